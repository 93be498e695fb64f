"""Drop-in for mlx_audio/codec/models/bigvgan/conv.py: ``WNConv1d`` (:15-62) and
``WNConvTranspose1d`` (:65-114) on the implicit-GEMM convolution kernels (3xTF32 on tcgen05 where shapes allow), through
the layer machinery of ``codec/models/snac/layers.py``.  Same constructor arguments, parameter names and MLX layouts:
both weights are (out, K, in); the transposed one is normalised per INPUT channel (``except_dim=2``)."""
from __future__ import annotations

import math

import torch

from .... import _lib as L
from .... import utils as U
from ..snac import layers as S

WNConv1d = S.WNConv1d      # conv.py:15-62: same weight norm (per output channel, no epsilon in the forward) and call


class WNConvTranspose1d(S.Module):
    """conv.py:65-114.  ``kernel_size = 2 * stride`` (every BigVGAN configuration: 8/4, 4/2, 16/8 ...) runs as the
    polyphase implicit GEMM of the Kokoro upsamplers.  Unlike SNAC / DAC this class passes its own ``output_padding``
    in the sixth position (:108-110), so L_out = (L - 1) * stride - 2 * padding + K + output_padding.
    ``in_scale``: the input is multiplied by it (folded into the packed weights) - BigVGAN divides the sum of a stage's
    resblocks by their number right before the next upsampler (bigvgan.py:117)."""

    def __init__(self, in_channels: int, out_channels: int, kernel_size: int, stride: int = 1, padding: int = 0,
                 dilation: int = 1, output_padding: int = 0, bias: bool = True):
        if kernel_size != 2 * stride or dilation != 1:
            raise NotImplementedError("bigvgan.WNConvTranspose1d: kernel_size = 2 * stride, dilation 1 only")
        self.kernel_size, self.stride, self.padding, self.dilation = kernel_size, stride, padding, dilation
        self.output_padding = output_padding
        scale = math.sqrt(1 / (in_channels * kernel_size))
        w = (torch.rand(out_channels, kernel_size, in_channels, generator=torch.Generator().manual_seed(0)) * 2 - 1) * scale
        self.weight_g = w.pow(2).sum(dim=(0, 1), keepdim=True).sqrt()          # (1, 1, in): except_dim = 2
        self.weight_v = w / (self.weight_g + 1e-12)
        self.bias = torch.zeros(out_channels) if bias else None
        self._folded = None

    def __call__(self, x):
        L.require_cuda()
        x = U._as_cuda(x).contiguous()
        cout, K, cin = self.weight_v.shape
        if self._folded is None:
            # (out, K, in) -> (in, K, out): the norm then runs over the rows of the leading axis, and it is the layout the
            # polyphase weight pack reads (kk_pack_conv_weight mode 1)
            v = U._as_cuda(self.weight_v).permute(2, 1, 0).contiguous()
            w = S._fold(v, U._as_cuda(self.weight_g).reshape(cin, 1, 1).contiguous())
            self._folded = dict(S._pack(w, 1, self.stride, cout % 32 == 0),
                                bias=None if self.bias is None else U._as_cuda(self.bias).contiguous())
        f = self._folded
        B, Lin, _ = x.shape
        s, p = self.stride, self.padding
        L_true = (Lin - 1) * s - 2 * p + K + self.output_padding
        return S._gemm_conv(x, f, taps=2, dil=1, pad_left=1, M=Lin + 1, N=s * cout, C_out=cout, up=s, up_off=p,
                            L_true=L_true, bias=f["bias"])
