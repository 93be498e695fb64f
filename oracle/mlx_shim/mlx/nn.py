"""``mlx.nn`` stand-in: just enough Module machinery for the reference's istftnet.py.

TEST INFRASTRUCTURE ONLY - see oracle/mlx_shim/mlx/__init__.py.
"""
from __future__ import annotations

import math

import numpy as np

from . import core as mx


class Module:
    training = False

    def __init__(self, *a, **k):
        pass

    def eval(self):
        return self

    def __contains__(self, key):      # mlx.nn.Module is a dict of its members ("bias" in self, descript/nn/layers.py:59)
        return vars(self).get(key) is not None

    # -- flat parameter access by the dotted names MLX would produce
    def named_parameters(self, prefix=""):
        out = {}

        def visit(obj, pfx):
            if isinstance(obj, np.ndarray):
                out[pfx] = obj
            elif isinstance(obj, Module):
                for k, v in vars(obj).items():
                    if not k.startswith("_"):
                        visit(v, f"{pfx}.{k}" if pfx else k)
            elif isinstance(obj, (list, tuple)):
                for i, v in enumerate(obj):
                    visit(v, f"{pfx}.{i}")

        visit(self, prefix)
        return out

    def set_parameter(self, dotted: str, value):
        obj = self
        parts = dotted.split(".")
        for p in parts[:-1]:
            obj = obj[int(p)] if isinstance(obj, (list, tuple)) else getattr(obj, p)
        last = parts[-1]
        value = mx.array(value)
        if isinstance(obj, list):
            cur = obj[int(last)]
            assert tuple(cur.shape) == tuple(value.shape), (dotted, cur.shape, value.shape)
            obj[int(last)] = value
        else:
            cur = getattr(obj, last)
            assert cur is not None and tuple(cur.shape) == tuple(value.shape), (dotted, getattr(cur, "shape", None), value.shape)
            setattr(obj, last, value)


class Linear(Module):
    def __init__(self, input_dims, output_dims, bias=True):
        scale = math.sqrt(1.0 / input_dims)
        self.weight = mx.array(np.random.uniform(-scale, scale, (output_dims, input_dims)))
        if bias:
            self.bias = mx.array(np.random.uniform(-scale, scale, (output_dims,)))

    def __call__(self, x):
        y = x @ self.weight.T
        if getattr(self, "bias", None) is not None:
            y = y + self.bias
        return y


class Conv1d(Module):
    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1, bias=True):
        self.weight = mx.zeros((int(out_channels), int(kernel_size), int(in_channels) // groups))
        if bias:
            self.bias = mx.zeros((int(out_channels),))
        self.stride, self.padding, self.dilation, self.groups = int(stride), int(padding), int(dilation), groups

    def __call__(self, x):
        y = mx.conv1d(x, self.weight, self.stride, self.padding, self.dilation, self.groups)
        if getattr(self, "bias", None) is not None:
            y = y + self.bias
        return y


class Upsample(Module):
    """Nearest-neighbour, integer scale, channels-last (N, L, C): element repeat along L."""

    def __init__(self, scale_factor, mode="nearest", align_corners=False):
        self._scale = int(scale_factor)
        assert mode == "nearest"

    def __call__(self, x):
        return mx.repeat(x, self._scale, axis=1)


class LeakyReLU(Module):
    def __init__(self, negative_slope=0.01):
        self._slope = negative_slope

    def __call__(self, x):
        return mx.maximum(self._slope * x, x)  # mlx.nn.leaky_relu


class Dropout(Module):
    def __init__(self, p=0.5):
        self._p = p

    def __call__(self, x):
        return x


class Identity(Module):
    def __call__(self, x, *a, **k):
        return x


class Sequential(Module):
    """mlx.nn.Sequential: members live in ``.layers`` (parameter names ``layers.<i>....``)."""

    def __init__(self, *modules):
        self.layers = list(modules)

    def __call__(self, x):
        for m in self.layers:
            x = m(x)
        return x


class Tanh(Module):
    def __call__(self, x):
        return mx.array(np.tanh(np.asarray(x)))


class ConvTranspose1d(Module):
    """Named (never instantiated) by codec/models/bigvgan/conv.py:111, a bare expression statement."""
