"""Drop-in for the Snake pieces of mlx_audio/codec/models/descript/nn/layers.py:116-129: ``snake``
and ``Snake1d`` (channels-last, alpha (1, 1, C)) on ``kk_snake_f32``.  The weight-normed convolutions of that file
(WNConv1d / WNConvTranspose1d, :15-113) normalise without the epsilon of Kokoro's ``weight_norm`` and take ``groups``:
not mapped onto the decoder's convolution kernels yet."""
from __future__ import annotations

import torch

from ...bigvgan.activation import _snake


def snake(x, alpha):
    """layers.py:116-119: x (B, T, C), alpha broadcastable (1, 1, C)."""
    return _snake(x, torch.as_tensor(alpha).reshape(-1), None, False, False)


class Snake1d:
    def __init__(self, channels: int):
        self.alpha = torch.ones((1, 1, channels))

    def __call__(self, x):
        return snake(x, self.alpha)
