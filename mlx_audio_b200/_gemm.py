"""y = x W^T + b on the tensor cores in fp32 accuracy: ``kk_conv1d_tf32_tc`` (tcgen05 ``kind::tf32`` with the 3xTF32 hi/lo
split) used as a 1x1 convolution.  The dense layers of the widened surface that are not convolutions go through this -
the LSTM input projection of ``ProsodyPredictor.shared`` (modules.py:93-286), Vocos' ``ISTFTHead.out`` (vocos.py:124) -
so that no library GEMM is left on those paths."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib as L


class LinearTF32x3:
    def __init__(self, weight: torch.Tensor, bias: Optional[torch.Tensor]):
        """weight (out, in) fp32 on the device; ``in`` must be a multiple of 4 (16-byte rows of the activation)."""
        from .tts.models.kokoro.engine import PreparedConv
        if weight.dim() != 2 or weight.shape[1] % 4:
            raise ValueError(f"LinearTF32x3: weight must be (out, in) with in % 4 == 0, got {tuple(weight.shape)}")
        self.n_out, self.n_in = int(weight.shape[0]), int(weight.shape[1])
        self.pc = PreparedConv("linear", weight[:, None, :].contiguous(), None if bias is None else bias.contiguous(),
                               need_bf16=False, need_f32=False, need_tf32=True)

    def __call__(self, x: torch.Tensor, out_pitch: Optional[int] = None) -> torch.Tensor:
        """x (B, L, in) fp32 contiguous -> (B, L, out_pitch) with the result in [..., :out] (out_pitch >= out, a multiple
        of 4 floats; default: out rounded up)."""
        if x.dim() != 3 or x.shape[2] != self.n_in or not x.is_contiguous() or x.dtype != torch.float32:
            raise ValueError(f"LinearTF32x3: x must be contiguous fp32 (B, L, {self.n_in}), got {tuple(x.shape)} {x.dtype}")
        B, L_, _ = x.shape
        pitch = out_pitch or -(-self.n_out // 4) * 4
        if pitch < self.n_out or pitch % 4:
            raise ValueError(f"LinearTF32x3: out_pitch {pitch} must be >= {self.n_out} and a multiple of 4")
        pc = self.pc
        y = torch.empty(B, L_, pitch, device=x.device, dtype=torch.float32)
        d = L.ConvDesc()
        d.B, d.L_in, d.C_in, d.x_pitch = B, L_, self.n_in, self.n_in
        d.taps, d.dil, d.pad_left, d.stride, d.Cw = 1, 1, 0, 1, pc.CwT
        d.M, d.N, d.C_out, d.up, d.up_off = L_, pc.N, pc.C_out, 1, 0
        d.L_true = d.L_out = L_
        d.bias, d.scale = (L.ptr(pc.bias) if pc.bias is not None else None), 1.0
        d.out, d.out_pitch = L.ptr(y), pitch
        L.check(L.load().kk_conv1d_tf32_tc(L.ptr(x), L.ptr(pc.w32hi), L.ptr(pc.w32lo), C.byref(d), 3, L.stream()),
                "kk_conv1d_tf32_tc")
        return y
