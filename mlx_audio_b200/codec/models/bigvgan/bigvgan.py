"""B200 drop-in for the BigVGAN generator of mlx_audio/codec/models/bigvgan/bigvgan.py (VERDICT r01
named BigVGAN / DAC / SNAC as the generators that reuse the decoder's conv + Snake kernels): ``BigVGANConfig`` (:14-26),
``BigVGAN`` (:29-128) and its ``sanitize`` (:130-149) - same configuration fields, parameter names, MLX layouts and call
contract ((B, num_mels, T) -> (B, 1, T * prod(upsample_rates))).

Kernels: the weight-normed convolutions are the implicit-GEMM kernels of the Kokoro decoder (3xTF32 on tcgen05 where the
channel counts allow, the polyphase form for the transposed ones); every ``Activation1d`` - two per AMP conv pair, the
costliest part of the reference (a depthwise transposed conv, a Snake and a depthwise strided conv over a 2x-rate signal)
- is the single fused ``kk_aa_act_f32``; the sum over a stage's resblocks and its division are ``kk_axpby_f32``; the
final ``tanh`` is ``kk_tanh_f32``.  ``activation="snake"`` is rejected: the reference's ``Snake`` broadcasts alpha over
the time axis of the channels-last tensors it is given here (activation.py:16-18) and only evaluates when T equals the
channel count.  No CPU path.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List

import numpy as np
import torch

from .... import _lib as L
from .... import utils as U
from ..snac.layers import Module
from .activation import Snake, SnakeBeta
from .amp import AMPBlock1, AMPBlock2
from .conv import WNConv1d, WNConvTranspose1d
from .resample import Activation1d


@dataclass
class BigVGANConfig:
    num_mels: int
    upsample_rates: List[int]
    upsample_kernel_sizes: List[int]
    upsample_initial_channel: int
    resblock: str                       # "1" | "2"
    resblock_kernel_sizes: List[int]
    resblock_dilation_sizes: List[List[int]]
    activation: str                     # "snakebeta" | "snake"
    snake_logscale: bool
    use_bias_at_final: bool = True
    use_tanh_at_final: bool = True


def _axpby(x, y, a: float, b: float):
    out = torch.empty_like(x)
    L.check(L.load().kk_axpby_f32(L.ptr(x), None if y is None else L.ptr(y), L.ptr(out), float(a), float(b), x.numel(),
                                  L.stream()), "kk_axpby_f32")
    return out


class BigVGAN(Module):
    def __init__(self, config: BigVGANConfig):
        c = config
        self.num_kernels = len(c.resblock_kernel_sizes)
        self.num_upsamples = len(c.upsample_rates)
        self.use_tanh_at_final = c.use_tanh_at_final
        C0 = c.upsample_initial_channel
        self.conv_pre = WNConv1d(c.num_mels, C0, 7, 1, 3)
        self.ups = [[WNConvTranspose1d(C0 // (2 ** i), C0 // (2 ** (i + 1)), k, u, padding=(k - u) // 2)]
                    for i, (u, k) in enumerate(zip(c.upsample_rates, c.upsample_kernel_sizes))]
        block = AMPBlock1 if c.resblock == "1" else AMPBlock2
        self.resblocks = [block(C0 // (2 ** (i + 1)), c.snake_logscale, c.activation, k, d)
                          for i in range(len(self.ups))
                          for k, d in zip(c.resblock_kernel_sizes, c.resblock_dilation_sizes)]
        Cl = C0 // (2 ** len(self.ups))
        self.activation_post = Activation1d(Snake(Cl, alpha_logscale=c.snake_logscale) if c.activation == "snake"
                                            else SnakeBeta(Cl, alpha_logscale=c.snake_logscale))
        self.conv_post = WNConv1d(Cl, 1, 7, 1, padding=3, bias=c.use_bias_at_final)

    def __call__(self, x, *args, **kwargs) -> torch.Tensor:
        """(batch, num_mels, seq) -> (batch, 1, seq * prod(upsample_rates))   bigvgan.py:102-128"""
        L.require_cuda()
        x = U._as_cuda(x)
        if x.dim() != 3:
            raise ValueError(f"BigVGAN expects (B, num_mels, T), got {tuple(x.shape)}")
        x = x.transpose(1, 2).contiguous()
        x = self.conv_pre(x)
        for step in range(self.num_upsamples):
            for up in self.ups[step]:
                x = up(x)
            xs = self.resblocks[step * self.num_kernels](x)
            for idx in range(1, self.num_kernels):
                xs = _axpby(xs, self.resblocks[step * self.num_kernels + idx](x), 1.0, 1.0)
            x = _axpby(xs, None, 1.0 / self.num_kernels, 0.0)
        x = self.activation_post(x)
        x = self.conv_post(x)
        if self.use_tanh_at_final:
            y = torch.empty_like(x)
            L.check(L.load().kk_tanh_f32(L.ptr(x), L.ptr(y), x.numel(), L.stream()), "kk_tanh_f32")
        else:
            y = _clip(x)
        return y.transpose(1, 2)

    def sanitize(self, weights):
        """bigvgan.py:130-149: drop ``num_batches_tracked``; transpose torch-layout conv / filter weights (out, in, K) ->
        (out, K, in) and transposed-conv weights (in, out, K) -> (out, K, in) when the shape differs from this model's."""
        cur = {k: tuple(v.shape) for k, v in self.named_parameters().items()}
        out = {}
        for key, value in (weights.items() if isinstance(weights, dict) else weights):
            if "num_batches_tracked" in key:
                continue
            value = np.asarray(value) if not isinstance(value, torch.Tensor) else value
            if "conv" in key or "lowpass.filter" in key or "upsample.filter" in key:
                if value.ndim == 3 and tuple(value.shape) != cur.get(key):
                    value = value.transpose(0, 2, 1) if isinstance(value, np.ndarray) else value.permute(0, 2, 1)
            if "ups." in key:
                if value.ndim == 3 and tuple(value.shape) != cur.get(key):
                    value = value.transpose(1, 2, 0) if isinstance(value, np.ndarray) else value.permute(1, 2, 0)
            out[key] = value
        return out


def _clip(x):
    """mx.clip(x, -1, 1) (:125-126)."""
    out = torch.empty_like(x)
    L.check(L.load().kk_clip_f32(L.ptr(x), L.ptr(out), -1.0, 1.0, x.numel(), L.stream()), "kk_clip_f32")
    return out
