"""Drop-in for the Snake pieces of mlx_audio/codec/models/descript/nn/layers.py:116-129: ``snake``
and ``Snake1d`` (channels-last, alpha (1, 1, C)) on ``kk_snake_f32``, and ``WNConv1d`` (:15-59) on the decoder's
weight-norm fold (``kk_fold_weight_norm_ex``: no epsilon, (g v) / ||v||) and fp32 implicit-GEMM convolution
(``kk_conv1d_f32``: stride, padding, dilation; ``groups`` = 1).  ``WNConvTranspose1d`` (:62-113, norm over another axis,
general stride / kernel) is not mapped yet."""
from __future__ import annotations

import ctypes as C

import torch

from ..... import _lib as L
from ..... import utils as U
from ...bigvgan.activation import _snake


class WNConv1d:
    """layers.py:15-59: ``weight_g`` (O, 1, 1), ``weight_v`` (O, K, I), ``bias`` (O); x (B, L, I) -> (B, L_out, O)."""

    def __init__(self, in_channels: int, out_channels: int, kernel_size: int, stride: int = 1, padding: int = 0,
                 dilation: int = 1, bias: bool = True, groups: int = 1):
        if groups != 1:
            raise NotImplementedError("WNConv1d: groups > 1 is not mapped onto the convolution kernels")
        self.kernel_size, self.stride, self.padding, self.dilation, self.groups = kernel_size, stride, padding, dilation, 1
        g = torch.Generator().manual_seed(0)
        scale = (1.0 / (in_channels * kernel_size)) ** 0.5
        w = (torch.rand(out_channels, kernel_size, in_channels, generator=g) * 2 - 1) * scale     # :38-43
        self.weight_g = w.pow(2).sum(dim=(1, 2), keepdim=True).sqrt()
        self.weight_v = w / (self.weight_g + 1e-12)
        self.bias = torch.zeros(out_channels) if bias else None

    def __call__(self, x):
        L.require_cuda()
        lib = L.load()
        x = U._as_cuda(x).contiguous()
        v, gq = U._as_cuda(self.weight_v).contiguous(), U._as_cuda(self.weight_g).contiguous()
        O, K, I = v.shape
        if x.dim() != 3 or x.shape[2] != I:
            raise ValueError(f"x must be (B, L, {I}), got {tuple(x.shape)}")
        B, Lin, _ = x.shape
        Lout = (Lin + 2 * self.padding - self.dilation * (K - 1) - 1) // self.stride + 1
        if Lout <= 0:
            raise ValueError("input too short for this kernel / dilation")
        w = torch.empty_like(v)
        L.check(lib.kk_fold_weight_norm_ex(L.ptr(v), L.ptr(gq), L.ptr(w), O, K * I, 0.0, 1, L.stream()),
                "kk_fold_weight_norm_ex")                                                         # :57
        bias = None if self.bias is None else U._as_cuda(self.bias).contiguous()
        out = torch.empty(B, Lout, O, dtype=torch.float32, device=x.device)
        d = L.ConvDesc()
        d.B, d.L_in, d.C_in, d.x_pitch = B, Lin, I, I
        d.taps, d.dil, d.pad_left, d.stride, d.Cw = K, self.dilation, self.padding, self.stride, I
        d.M, d.N, d.C_out, d.up, d.up_off = Lout, O, O, 1, 0
        d.L_true = d.L_out = Lout
        d.left_pad = 0
        d.bias = None if bias is None else bias.data_ptr()
        d.scale = 1.0
        d.out, d.out_pitch = out.data_ptr(), O
        L.check(lib.kk_conv1d_f32(L.ptr(x), L.ptr(w), C.byref(d), L.stream()), "kk_conv1d_f32")    # :58-61
        self._keep = (x, w, bias, d)
        return out


def snake(x, alpha):
    """layers.py:116-119: x (B, T, C), alpha broadcastable (1, 1, C)."""
    return _snake(x, torch.as_tensor(alpha).reshape(-1), None, False, False)


class Snake1d:
    def __init__(self, channels: int):
        self.alpha = torch.ones((1, 1, channels))

    def __call__(self, x):
        return snake(x, self.alpha)
