"""Drop-in for the ``ISTFTHead`` of mlx_audio/codec/models/vocos/vocos.py:119-140: the same iSTFT
kernels as ``utils.istft`` behind a Linear projection.  The projection runs on the tensor cores in fp32 accuracy
(``_gemm.LinearTF32x3``: ``kk_conv1d_tf32_tc`` as a 1x1 convolution; it was a cuBLAS call in round 1); the
polar stage (exp, clip, cos, sin) is ``kk_polar_to_spec`` writing the frame-major spectrum ``kk_istft_f32`` reads, so
no transposed copy is made.  No CPU path."""
from __future__ import annotations

import numpy as np
import torch

from .... import _lib as L
from .... import utils as U


class ISTFTHead:
    def __init__(self, dim: int, n_fft: int, hop_length: int, padding: str = "center", device=None):
        self.n_fft, self.hop_length, self.dim = n_fft, hop_length, dim
        self.device = torch.device(device or "cuda")
        self.weight = None        # out.weight (n_fft + 2, dim)
        self.bias = None          # out.bias   (n_fft + 2,)
        self._lin = None

    def load_weights(self, weights, strict: bool = True):
        """``out.weight`` / ``out.bias`` (with or without a ``head.`` prefix)."""
        L.require_cuda()
        items = dict(weights.items() if isinstance(weights, dict) else weights)
        for k, v in items.items():
            k = k[len("head."):] if k.startswith("head.") else k
            t = torch.as_tensor(np.asarray(v) if not isinstance(v, torch.Tensor) else v, dtype=torch.float32)
            if k == "out.weight":
                if tuple(t.shape) != (self.n_fft + 2, self.dim):
                    raise ValueError(f"out.weight must be {(self.n_fft + 2, self.dim)}, got {tuple(t.shape)}")
                self.weight = t.to(self.device).contiguous()
                self._lin = None
            elif k == "out.bias":
                if tuple(t.shape) != (self.n_fft + 2,):
                    raise ValueError(f"out.bias must be {(self.n_fft + 2,)}, got {tuple(t.shape)}")
                self.bias = t.to(self.device).contiguous()
                self._lin = None
            elif strict:
                raise ValueError(f"Unexpected ISTFTHead parameter {k!r}")
        return self

    def __call__(self, x) -> torch.Tensor:
        """x (1, L, dim) -> audio ((L - 1) * hop,) like the reference (:125-140); extension: (B, L, dim) -> (B, len)."""
        L.require_cuda()
        if self.weight is None or self.bias is None:
            raise L.KokoroB200Error("ISTFTHead weights are not loaded")
        x = U._as_cuda(x)
        if x.dim() != 3 or x.shape[2] != self.dim:
            raise ValueError(f"x must be (B, L, {self.dim}), got {tuple(x.shape)}")
        B, frames, _ = x.shape
        bins = self.n_fft // 2 + 1
        if self._lin is None:
            from ...._gemm import LinearTF32x3
            self._lin = LinearTF32x3(self.weight, self.bias)
        pitch = -(-(self.n_fft + 2) // 4) * 4
        h = self._lin(x.contiguous(), pitch)                                                # nn.Linear :124
        spec = torch.empty(B, frames, bins, 2, dtype=torch.float32, device=x.device)
        L.check(L.load().kk_polar_to_spec(L.ptr(h), pitch, L.ptr(spec), B * frames, bins, 1e2, L.stream()),
                "kk_polar_to_spec")                                                        # :126-133
        # istft(S, window=hanning(n_fft) [symmetric, explicit array], hop_length, win_length=n_fft) :134-139
        y = U.istft(torch.view_as_complex(spec).transpose(1, 2), hop_length=self.hop_length, win_length=self.n_fft,
                    window=U.hanning(self.n_fft))
        return y[0] if B == 1 else y
