"""Drop-in for mlx_audio/codec/models/bigvgan/activation.py: ``Snake`` (channels-first,
alpha[None, :, None]) and ``SnakeBeta`` (channels-last, separate beta) on ``kk_snake_f32``.  Parameters keep the
reference's names (``alpha``, ``beta``) and initial values."""
from __future__ import annotations

import torch

from .... import _lib as L
from .... import utils as U


def _snake(x, alpha, beta, channels_first: bool, logscale: bool):
    L.require_cuda()
    x = U._as_cuda(x).contiguous()
    if x.dim() != 3:
        raise ValueError(f"expected a 3-D tensor, got {tuple(x.shape)}")
    C = x.shape[1] if channels_first else x.shape[2]
    alpha = U._as_cuda(alpha).contiguous()
    beta = None if beta is None else U._as_cuda(beta).contiguous()
    if alpha.numel() != C or (beta is not None and beta.numel() != C):
        raise ValueError(f"alpha / beta must have {C} entries")
    out = torch.empty_like(x)
    outer = x.shape[0] if channels_first else x.shape[0] * x.shape[1]
    inner = x.shape[2] if channels_first else 1
    L.check(L.load().kk_snake_f32(L.ptr(x), L.ptr(out), L.ptr(alpha), None if beta is None else L.ptr(beta), outer, C,
                                  inner, int(logscale), 1e-9, L.stream()), "kk_snake_f32")
    return out


class Snake:
    """activation.py:5-24: x (B, C, T)."""

    def __init__(self, in_features: int, alpha: float = 1.0, alpha_logscale: bool = False):
        self.alpha_logscale = alpha_logscale
        self.alpha = (torch.zeros(in_features) if alpha_logscale else torch.ones(in_features)) * alpha

    def __call__(self, x):
        return _snake(x, self.alpha, None, True, self.alpha_logscale)


class SnakeBeta:
    """activation.py:27-51: x (B, T, C)."""

    def __init__(self, in_features: int, alpha: float = 1.0, alpha_logscale: bool = False):
        self.alpha_logscale = alpha_logscale
        self.alpha = (torch.zeros(in_features) if alpha_logscale else torch.ones(in_features)) * alpha
        self.beta = (torch.zeros(in_features) if alpha_logscale else torch.ones(in_features)) * alpha

    def __call__(self, x):
        return _snake(x, self.alpha, self.beta, False, self.alpha_logscale)
