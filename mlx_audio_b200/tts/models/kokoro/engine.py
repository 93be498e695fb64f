"""Launch plan of the Kokoro ISTFTNet decoder on libkokoro_b200 (host side, Python + torch plumbing only).

The reference executes ``Decoder.__call__`` (mlx_audio/tts/models/kokoro/istftnet.py:947-963) as ~600 lazy MLX
ops on channels-first tensors.  Here the same arithmetic is laid out once per (batch, frames) shape as a
flat list of C-ABI kernel launches over persistent channels-last buffers:

* "pre" tensors  (fp32, (B, L, C)): conv outputs / residual streams, written by conv epilogues together with
  their per-(b, c) instance-norm sums;
* "act" tensors  (bf16 in tensor-core mode, fp32 in strict mode, (B, L, Cp)): AdaIN + Snake / LeakyReLU
  outputs, the A operands of the next implicit-GEMM convolution.

No torch op runs on the hot path except one ``zero_()`` of the statistics arena.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Callable, Dict, List, Optional, Tuple

import torch

from ... import _lib as L
from ..._lib import ConvDesc


def _ceil(a: int, m: int) -> int:
    return (a + m - 1) // m * m


class PreparedConv:
    """A conv layer folded (weight norm) and packed for both convolution kernels."""

    def __init__(self, name: str, w_folded: torch.Tensor, bias: Optional[torch.Tensor], transposed_up: int = 0,
                 need_bf16: bool = True, need_f32: bool = True):
        lib = L.load()
        st = L.stream()
        self.name = name
        O, K, I = w_folded.shape
        self.up = transposed_up if transposed_up else 1
        if transposed_up:
            self.C_in, self.C_out, self.taps = O, I, 2
            self.N = transposed_up * I
            mode = 1
        else:
            self.C_in, self.C_out, self.taps = I, O, K
            self.N = O
            mode = 0
        self.bias = bias
        self.Cw32 = self.C_in
        self.Cw16 = _ceil(self.C_in, 64)
        dev = w_folded.device
        self.w32 = self.w16 = None
        if need_f32:
            if mode == 0:
                self.w32 = w_folded.contiguous()
            else:
                self.w32 = torch.empty(self.N, self.taps * self.Cw32, dtype=torch.float32, device=dev)
                L.check(lib.kk_pack_conv_weight(L.ptr(w_folded), L.ptr(self.w32), 0, mode, O, K, I, self.up, self.Cw32,
                                                st), "kk_pack_conv_weight")
        if need_bf16:
            self.w16 = torch.empty(self.N, self.taps * self.Cw16, dtype=torch.bfloat16, device=dev)
            L.check(lib.kk_pack_conv_weight(L.ptr(w_folded), L.ptr(self.w16), 1, mode, O, K, I, self.up, self.Cw16, st),
                    "kk_pack_conv_weight")


class Plan:
    """All buffers and the launch list for one (B, T) shape."""

    def __init__(self):
        self.ops: List[Tuple[Callable, tuple]] = []
        self.keep: List[object] = []          # ctypes structs / tensors that must outlive the plan
        self.buffers: Dict[str, torch.Tensor] = {}
        self.taps: Dict[str, Tuple[torch.Tensor, tuple]] = {}
        self.n_launches = 0
        self.stats_arena: Optional[torch.Tensor] = None

    def add(self, fn, *args, launches: int = 1):
        self.ops.append((fn, args))
        self.n_launches += launches

    def run(self, stream: int):
        for fn, args in self.ops:
            rc = fn(*args, stream)
            if rc != 0:
                L.check(rc, getattr(fn, "__name__", "kernel"))
