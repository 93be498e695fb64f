"""Drop-in for mlx_audio/codec/models/bigvgan/resample.py: the kaiser-sinc filter design
(:18-50) and ``Activation1d`` (:152-177) - upsample x2, activation, low-pass + downsample x2 - as ONE kernel
(``kk_aa_act_f32``): the 2x-rate signal never exists in memory.  Channels-last (B, T, C) like the reference.

Only the configuration BigVGAN builds is supported (ratios 2 / 2, 12-tap filters: the ``Activation1d`` defaults the AMP
blocks and ``activation_post`` use, amp.py:43-50, bigvgan.py:79-88).  ``upsample.filter`` and
``downsample.lowpass.filter`` are parameters with the reference's names and (1, K, 1) shape, so a checkpoint's copies
are loaded over the designed ones.
"""
from __future__ import annotations

import math

import numpy as np
import torch

from .... import _lib as L
from .... import utils as U
from ..snac.layers import Module
from .activation import Snake, SnakeBeta


def kaiser_sinc_filter1d(cutoff: float, half_width: float, kernel_size: int) -> torch.Tensor:
    """resample.py:18-50, same operation order in fp32; returns (1, kernel_size, 1)."""
    even = kernel_size % 2 == 0
    half_size = kernel_size // 2
    delta_f = 4 * half_width
    A = 2.285 * (half_size - 1) * math.pi * delta_f + 7.95
    if A > 50.0:
        beta = 0.1102 * (A - 8.7)
    elif A >= 21.0:
        beta = 0.5842 * (A - 21) ** 0.4 + 0.07886 * (A - 21.0)
    else:
        beta = 0.0
    window = np.kaiser(kernel_size, beta).astype(np.float32)            # mx.array(np.kaiser(...)): float64 -> float32
    if even:
        time = np.arange(-half_size, half_size).astype(np.float32) + np.float32(0.5)
    else:
        time = (np.arange(kernel_size) - half_size).astype(np.float32)
    if cutoff == 0:
        return torch.zeros(1, kernel_size, 1)
    t = (np.float32(2 * cutoff) * time).astype(np.float32)
    with np.errstate(invalid="ignore", divide="ignore"):
        sinc = np.where(t == 0, np.float32(1.0), np.sin(np.float32(math.pi) * t) / np.float32(math.pi) / t).astype(np.float32)
    filt = (np.float32(2 * cutoff) * window * sinc).astype(np.float32)
    filt = (filt / filt.sum(dtype=np.float32)).astype(np.float32)
    return torch.from_numpy(filt.reshape(1, kernel_size, 1))


class LowPassFilter1d(Module):
    """Holder of ``filter`` (resample.py:53-101); the convolution itself is inside the fused kernel."""

    def __init__(self, cutoff: float = 0.5, half_width: float = 0.6, stride: int = 1, kernel_size: int = 12):
        if cutoff < -0.0:
            raise ValueError("Minimum cutoff must be larger than zero.")
        if cutoff > 0.5:
            raise ValueError("A cutoff above 0.5 does not make sense.")
        self.stride, self.kernel_size = stride, kernel_size
        self.filter = kaiser_sinc_filter1d(cutoff, half_width, kernel_size)


class UpSample1d(Module):
    def __init__(self, ratio: int = 2, kernel_size=None):
        self.ratio = ratio
        self.kernel_size = int(6 * ratio // 2) * 2 if kernel_size is None else kernel_size
        self.filter = kaiser_sinc_filter1d(0.5 / ratio, 0.6 / ratio, self.kernel_size)


class DownSample1d(Module):
    def __init__(self, ratio: int = 2, kernel_size=None):
        self.ratio = ratio
        self.kernel_size = int(6 * ratio // 2) * 2 if kernel_size is None else kernel_size
        self.lowpass = LowPassFilter1d(0.5 / ratio, 0.6 / ratio, ratio, self.kernel_size)


class _SnakeBetaModule(SnakeBeta, Module):
    """``SnakeBeta`` with parameter discovery (``act.alpha`` / ``act.beta`` in ``named_parameters``)."""


class Activation1d(Module):
    def __init__(self, activation, up_ratio: int = 2, down_ratio: int = 2, up_kernel_size: int = 12,
                 down_kernel_size: int = 12):
        if (up_ratio, down_ratio, up_kernel_size, down_kernel_size) != (2, 2, 12, 12):
            raise NotImplementedError("Activation1d: ratios 2 / 2 with 12-tap filters only (what BigVGAN builds)")
        if isinstance(activation, Snake):
            # activation.py:16-18 broadcasts alpha as [None, :, None] - over the TIME axis of this channels-last tensor;
            # in the reference that only evaluates when T equals the channel count.  Not reproduced.
            raise NotImplementedError("Activation1d(Snake): the reference's Snake is channels-first and does not broadcast "
                                      "on the (B, T, C) tensors BigVGAN feeds it; every released config uses snakebeta")
        if not isinstance(activation, SnakeBeta):
            raise NotImplementedError("Activation1d: SnakeBeta only")
        self.up_ratio, self.down_ratio = up_ratio, down_ratio
        if not isinstance(activation, Module):
            act = _SnakeBetaModule.__new__(_SnakeBetaModule)
            act.__dict__.update(activation.__dict__)
            activation = act
        self.act = activation
        self.upsample = UpSample1d(up_ratio, up_kernel_size)
        self.downsample = DownSample1d(down_ratio, down_kernel_size)
        self._host = None          # (filter tensors the copies were taken from, host copies of their 12 taps)

    def _filters(self):
        f, g = self.upsample.filter, self.downsample.lowpass.filter
        if self._host is None or self._host[0] is not f or self._host[1] is not g:     # load_weights replaces the tensors
            self._host = (f, g, f.detach().to("cpu", torch.float32).contiguous().reshape(-1),
                          g.detach().to("cpu", torch.float32).contiguous().reshape(-1))
        return self._host[2], self._host[3]

    def __call__(self, x):
        L.require_cuda()
        x = U._as_cuda(x).contiguous()
        if x.dim() != 3 or x.shape[2] != self.act.alpha.numel():
            raise ValueError(f"Activation1d expects (B, T, {self.act.alpha.numel()}), got {tuple(x.shape)}")
        B, T, C = x.shape
        out = torch.empty_like(x)
        alpha, beta = U._as_cuda(self.act.alpha).contiguous(), U._as_cuda(self.act.beta).contiguous()
        up, dn = self._filters()
        L.check(L.load().kk_aa_act_f32(L.ptr(x), L.ptr(out), L.ptr(alpha), L.ptr(beta), up.data_ptr(), dn.data_ptr(), B, T, C,
                                       int(self.act.alpha_logscale), 1e-9, L.stream()), "kk_aa_act_f32")
        return out
