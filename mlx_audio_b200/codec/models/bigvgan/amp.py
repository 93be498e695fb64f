"""Drop-in for mlx_audio/codec/models/bigvgan/amp.py: ``AMPBlock1`` (:10-60) and ``AMPBlock2``
(:63-96) - weight-normed dilated convolutions between anti-aliased periodic activations, each pair closed by a residual
add that runs in the convolution's epilogue."""
from __future__ import annotations

from typing import List

from .... import utils as U
from ..snac.layers import Module
from .activation import Snake, SnakeBeta
from .conv import WNConv1d
from .resample import Activation1d


def _activation(channels: int, snake_logscale: bool, activation: str) -> Activation1d:
    return Activation1d(Snake(channels, alpha_logscale=snake_logscale) if activation == "snake"
                        else SnakeBeta(channels, alpha_logscale=snake_logscale))


class AMPBlock1(Module):
    def __init__(self, channels: int, snake_logscale: bool, activation: str, kernel_size: int = 3,
                 dilation: List[int] = (1, 3, 5)):
        self.convs1 = [WNConv1d(channels, channels, kernel_size, stride=1, dilation=d, padding=((kernel_size - 1) * d) // 2)
                       for d in dilation]
        self.convs2 = [WNConv1d(channels, channels, kernel_size, stride=1, dilation=1, padding=(kernel_size - 1) // 2)
                       for _ in dilation]
        self.activations = [_activation(channels, snake_logscale, activation) for _ in range(len(dilation) * 2)]

    def __call__(self, x):
        x = U._as_cuda(x).contiguous()
        for c1, c2, a1, a2 in zip(self.convs1, self.convs2, self.activations[::2], self.activations[1::2]):
            x = c2(a2(c1(a1(x))), res=x)                     # x + conv2(act2(conv1(act1(x))))   amp.py:55-58
        return x


class AMPBlock2(Module):
    def __init__(self, channels: int, snake_logscale: bool, activation: str, kernel_size: int = 3,
                 dilation: List[int] = (1, 3, 5)):
        self.convs = [WNConv1d(channels, channels, kernel_size, stride=1, dilation=d, padding=((kernel_size - 1) * d) // 2)
                      for d in dilation]
        self.activations = [_activation(channels, snake_logscale, activation) for _ in dilation]

    def __call__(self, x):
        x = U._as_cuda(x).contiguous()
        for conv, act in zip(self.convs, self.activations):
            x = conv(act(x), res=x)                          # x + conv(act(x))   amp.py:92-94
        return x
