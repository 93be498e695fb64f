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
import os
from typing import Callable, Dict, List, Optional, Tuple

import torch

from .... import _lib as L
from ...._lib import AdainIn, ConvDesc


_NVTX = os.environ.get("KK_NVTX", "0") == "1"   # NVTX range per decoder stage (profiling aid, off by default)


def _ceil(a: int, m: int) -> int:
    return (a + m - 1) // m * m


class PreparedConv:
    """A conv layer folded (weight norm) and packed for both convolution kernels."""

    def __init__(self, name: str, w_folded: torch.Tensor, bias: Optional[torch.Tensor], transposed_up: int = 0,
                 need_bf16: bool = True, need_f32: bool = True, need_tf32: bool = False):
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
        self.w32hi = self.w32lo = None
        self.CwT = _ceil(self.C_in, 32)
        if need_tf32:
            # fp32 operands of the kind::tf32 kernel: K-major pack with 32-channel blocks, split into hi + lo (3xTF32)
            wt = torch.empty(self.N, self.taps * self.CwT, dtype=torch.float32, device=dev)
            L.check(lib.kk_pack_conv_weight(L.ptr(w_folded), L.ptr(wt), 0, mode, O, K, I, self.up, self.CwT, st),
                    "kk_pack_conv_weight")
            self.w32hi, self.w32lo = torch.empty_like(wt), torch.empty_like(wt)
            L.check(lib.kk_split_tf32(L.ptr(wt), L.ptr(self.w32hi), L.ptr(self.w32lo), wt.numel(), st), "kk_split_tf32")
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
        self.arenas: List[torch.Tensor] = []   # instance-norm (sum, sumsq) accumulators, zeroed every forward
        self._arena_used = 0
        self.captured: Dict[str, torch.Tensor] = {}
        self.capture = False
        self.zero_bufs: List[torch.Tensor] = []  # buffers whose pad rows / columns must read as zero
        self.conv_ops: Dict[int, ConvDesc] = {}  # op index -> descriptor (for FLOP accounting)
        # ragged batch (per-utterance frame counts in one launch, kokoro_b200.h): device address of the (B,) int32
        # lengths, and the row count of every level as (mul, add): rows = mul * frames + add
        self.lens_addr: Optional[int] = None
        self.levels: Dict[int, Tuple[int, int]] = {}
        # branch parallelism: tag of every op (0 = main path, 1 = the generator's noise branch, which depends on F0 and
        # the style projection only); fork_index = first op that may overlap the branch
        self.tags: List[int] = []
        self.cur_branch = 0
        self.fork_index: Optional[int] = None

    def _push(self, op):
        self.ops.append(op)
        self.tags.append(self.cur_branch)

    def add(self, fn, *args, launches: int = 1):
        self._push((fn, args))
        self.n_launches += launches

    def lens(self, rows: int) -> Tuple[Optional[int], int, int]:
        """(lens, len_mul, len_add) arguments of a kernel working on `rows` rows per (padded) utterance."""
        if self.lens_addr is None:
            return None, 0, 0
        mul, add = self.levels[rows]
        return self.lens_addr, mul, add

    def stage(self, name: str):
        """Stage boundary (front / source / stage0 / stage1 / post): an NVTX range when KK_NVTX=1 - the counterpart of the
        BenchmarkTimer spans of the reference's Swift port (Swift-TTS/Kokoro/Decoder.swift:74-105)."""
        self._push((None, ("@stage", name)))

    def _call(self, fn, args, stream: int):
        rc = fn(*args, stream)
        if rc != 0:
            L.check(rc, getattr(fn, "__name__", "kernel"))

    def run_parallel(self, side: "torch.cuda.Stream"):
        """The same launches with the noise branch (tag 1: harmonic source, its STFT, noise_convs, noise_res) on ``side``:
        forked after the style projection, joined before each upsampler (``@join`` markers wait for the ``@done`` events).
        Inside a CUDA-graph capture this becomes a parallel branch of the graph - at small batches, where no kernel fills
        the 148 SMs, it takes 17 of 97 launches off the critical path.  No debug taps, no NVTX ranges here."""
        main = torch.cuda.current_stream()
        ms, ss = main.cuda_stream, side.cuda_stream
        for a in self.arenas:
            a.zero_()
        n, fork = len(self.ops), self.fork_index
        for i in range(fork):
            fn, args = self.ops[i]
            if fn is not None:
                self._call(fn, args, ms)
        side.wait_event(main.record_event())
        done = {}
        with torch.cuda.stream(side):
            for i in range(fork, n):
                if self.tags[i] != 1:
                    continue
                fn, args = self.ops[i]
                if fn is not None:
                    self._call(fn, args, ss)
                elif args[0] == "@done":
                    done[args[1]] = side.record_event()
        for i in range(fork, n):
            if self.tags[i] != 0:
                continue
            fn, args = self.ops[i]
            if fn is not None:
                self._call(fn, args, ms)
            elif args[0] == "@join":
                main.wait_event(done[args[1]])

    def run(self, stream: int):
        nvtx = torch.cuda.nvtx if _NVTX else None
        open_range = False
        for a in self.arenas:
            a.zero_()
        for fn, args in self.ops:
            if fn is None:  # stage marker / debug tap
                if args[0] in ("@join", "@done"):
                    continue
                if args[0] == "@stage":
                    if nvtx is not None:
                        if open_range:
                            nvtx.range_pop()
                        nvtx.range_push("kokoro." + args[1])
                        open_range = True
                    continue
                if self.capture:
                    t, view = self.taps[args[0]]
                    self.captured[args[0]] = (view(t) if view else t).detach().float().clone()
                continue
            rc = fn(*args, stream)
            if rc != 0:
                L.check(rc, getattr(fn, "__name__", "kernel"))
        if open_range:
            nvtx.range_pop()

    def run_timed(self, stream: int):
        """One forward with a CUDA-event pair around every C-ABI call (on the launching stream).
        Returns [(symbol, op_index, milliseconds)].  Diagnostic only: the gaps serialise the launches."""
        for a in self.arenas:
            a.zero_()
        recs = []
        for i, (fn, args) in enumerate(self.ops):
            if fn is None:
                continue
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            rc = fn(*args, stream)
            e1.record()
            if rc != 0:
                L.check(rc, getattr(fn, "__name__", "kernel"))
            recs.append((fn.__name__, i, e0, e1))
        torch.cuda.synchronize()
        return [(n, i, e0.elapsed_time(e1)) for n, i, e0, e1 in recs]

    def conv_flops(self, op_index: int) -> float:
        d = self.conv_ops[op_index]
        return 2.0 * d.B * d.M * d.N * d.taps * d.C_in


class DecoderEngine:
    """Folds/packs the weights once, then builds and runs launch plans.

    ``weights``: flat dict with the reference's parameter names without the ``decoder.`` prefix
    (e.g. ``generator.resblocks.0.convs1.0.weight_v``), MLX layout (O, K, I), float32 CUDA tensors.
    ``mode``: "bf16" (tcgen05 tensor-core convolutions, bf16 operands / fp32 accumulate) or
    "fp32" (CUDA-core float32 convolutions, the strict parity mode).
    """

    def __init__(self, weights: Dict[str, torch.Tensor], cfg: dict, mode: str = "bf16", fused: bool = True,
                 res16: int = 1, fp32_conv: str = "tf32x3"):
        """``fused`` / ``res16`` select engine arms of the bf16 mode (defaults = the measured-fastest; every arm has a
        parity test, tests/test_gpu_decoder.py::test_decoder_bf16_engine_arms): fused = AdaIN + Snake inside the
        convolution's operand path (False: separate elementwise kernels, the layout the fp32 mode uses); res16 = bf16
        residual stream inside AdaINResBlock1 (0 off, 1 the 128-channel stage, 2 both stages)."""
        if mode not in ("bf16", "fp32"):
            raise ValueError(f"mode must be 'bf16' or 'fp32', got {mode!r}")
        if res16 not in (0, 1, 2):
            raise ValueError(f"res16 must be 0, 1 or 2, got {res16!r}")
        if fp32_conv not in ("tf32x3", "tf32", "ffma"):
            raise ValueError(f"fp32_conv must be 'tf32x3', 'tf32' or 'ffma', got {fp32_conv!r}")
        L.require_cuda()
        self.lib = L.load()
        self.cfg = cfg
        self.mode = mode
        self.tc = mode == "bf16"
        # fp32 mode: where the stride-1 convolutions run - "tf32x3" = tcgen05 kind::tf32 with the error-compensated
        # hi/lo split (fp32-grade, the default), "tf32" = plain TF32 (reported, misses the 1e-3 gate), "ffma" = CUDA cores
        self.fp32_conv = fp32_conv if (mode == "fp32" and self.lib.kk_device_is_sm100()) else "ffma"
        self.fused = self.tc and bool(fused)
        self.res16 = int(res16) if self.fused else 0
        if self.tc and not self.lib.kk_device_is_sm100():
            raise L.KokoroB200Error("bf16 tensor-core mode needs an sm_100 (B200) device")
        self.dev = next(iter(weights.values())).device
        self.w = weights
        self.convs: Dict[str, PreparedConv] = {}
        self.plans: Dict[Tuple[int, int], Plan] = {}
        # Workspace shared by every plan (one forward runs at a time, on one stream): a named pool of byte buffers,
        # each as large as the largest shape built so far.  Ragged workloads (one plan per (B, T) bucket) then cost
        # the memory of the largest plan, not the sum (161 distinct lengths).
        self.pool: Dict[str, torch.Tensor] = {}
        self.arena_pool: List[torch.Tensor] = []
        self.pool_grew = False
        self.active_plan: Optional[Plan] = None
        self.pools: Dict[str, Tuple[torch.Tensor, torch.Tensor]] = {}   # depthwise ConvTranspose "pool" per block
        self._check_cfg()
        self._prepare()

    def _check_cfg(self):
        cfg = self.cfg
        self.up_scale = int(math.prod(cfg["upsample_rates"])) * cfg["gen_istft_hop_size"]
        assert cfg["gen_istft_n_fft"] == 20 and cfg["gen_istft_hop_size"] == 5, "iSTFT head is specialised for 20/5"

    # ------------------------------------------------------------------ weights
    def _fold(self, prefix: str) -> torch.Tensor:
        v = self.w[f"{prefix}.weight_v"].contiguous()
        g = self.w[f"{prefix}.weight_g"].contiguous()
        out = torch.empty_like(v)
        O = v.shape[0]
        L.check(self.lib.kk_fold_weight_norm(L.ptr(v), L.ptr(g), L.ptr(out), O, v.numel() // O, L.stream()),
                "kk_fold_weight_norm")
        return out

    def _prep_wn(self, prefix: str, up: int = 0, tc: Optional[bool] = None):
        tc = self.tc if tc is None else tc
        self.convs[prefix] = PreparedConv(prefix, self._fold(prefix), self.w.get(f"{prefix}.bias"), up,
                                          need_bf16=tc, need_f32=not tc,
                                          need_tf32=(not tc) and self.fp32_conv != "ffma")

    def _prepare_adain(self):
        """AdaIN fc layers: one concatenated (R, style) matrix -> a single projection launch per forward."""
        w = self.w
        self.ada_off: Dict[str, int] = {}
        ws, bs, off = [], [], 0
        for k in sorted(k for k in w if k.endswith(".fc.weight")):
            p = k[: -len(".fc.weight")]
            self.ada_off[p] = off
            ws.append(w[k])
            bs.append(w[f"{p}.fc.bias"])
            off += w[k].shape[0]
        self.ada_W = torch.cat(ws, 0).contiguous()
        self.ada_b = torch.cat(bs, 0).contiguous()
        self.ada_R = off

    def _prep_resblk1d(self, blk: str):
        """Weights of one AdainResBlk1d (istftnet.py:853-861): conv1, conv2, [conv1x1], [pool]."""
        w = self.w
        self._prep_wn(f"{blk}.conv1")
        self._prep_wn(f"{blk}.conv2")
        if f"{blk}.conv1x1.weight_v" in w:
            self._prep_wn(f"{blk}.conv1x1")
        if f"{blk}.pool.weight_v" in w:
            self.pools[blk] = (self._fold(f"{blk}.pool").reshape(-1, 3).contiguous(), w[f"{blk}.pool.bias"].contiguous())

    def _prepare(self):
        w, cfg = self.w, self.cfg
        self._prepare_adain()
        # front
        for blk in ["encode", "decode.0", "decode.1", "decode.2", "decode.3"]:
            self._prep_resblk1d(blk)
        self._prep_wn("asr_res.0", tc=False)
        self.f0_w3b = torch.cat([self._fold("F0_conv").reshape(-1), w["F0_conv.bias"].reshape(-1)]).contiguous()
        self.n_w3b = torch.cat([self._fold("N_conv").reshape(-1), w["N_conv.bias"].reshape(-1)]).contiguous()
        # generator
        g = "generator"
        n_up = len(cfg["upsample_rates"])
        n_k = len(cfg["resblock_kernel_sizes"])
        for i, u in enumerate(cfg["upsample_rates"]):
            assert cfg["upsample_kernel_sizes"][i] == 2 * u, "polyphase path needs kernel == 2*stride"
            self._prep_wn(f"{g}.ups.{i}", up=u)
            nw, nbias = w[f"{g}.noise_convs.{i}.weight"].contiguous(), w[f"{g}.noise_convs.{i}.bias"]
            self.convs[f"{g}.noise_convs.{i}"] = PreparedConv(f"{g}.noise_convs.{i}", nw, nbias, need_bf16=False,
                                                              need_f32=not self.tc)
            if self.tc:
                # tensor-core form over the bf16 `har` tensor (B, rows, 24): a stride-s conv with kernel 2s and
                # padding p becomes a 3-tap stride-1 conv over rows of s frames (s*24 channels):
                #   frame = s*(m-1+t') + j  <->  tap t = s*t' + j - s + p   (zero weight when t is out of range)
                O, K, I = nw.shape
                s_ = int(math.prod(cfg["upsample_rates"][i + 1:]))
                if K == 1:
                    v = torch.zeros(O, 1, 24, device=nw.device)
                    v[:, 0, :I] = nw[:, 0]
                else:
                    assert K == 2 * s_, "tensor-core noise conv needs kernel == 2*stride"
                    p_ = (s_ + 1) // 2
                    v = torch.zeros(O, 3, s_, 24, device=nw.device)
                    for tp in range(3):
                        for j in range(s_):
                            t = s_ * tp + j - s_ + p_
                            if 0 <= t < K:
                                v[:, tp, j, :I] = nw[:, t]
                    v = v.reshape(O, 3, s_ * 24)
                self.convs[f"{g}.noise_convs.{i}.tc"] = PreparedConv(f"{g}.noise_convs.{i}.tc", v.contiguous(), nbias,
                                                                     need_bf16=True, need_f32=False)
            for blk in [f"{g}.noise_res.{i}"] + [f"{g}.resblocks.{i * n_k + j}" for j in range(n_k)]:
                for it in range(3):
                    self._prep_wn(f"{blk}.convs1.{it}")
                    self._prep_wn(f"{blk}.convs2.{it}")
        self._prep_wn(f"{g}.conv_post")
        self.lin_w = w[f"{g}.m_source.l_linear.weight"].reshape(-1).contiguous()
        self.lin_b = w[f"{g}.m_source.l_linear.bias"].reshape(-1).contiguous()
        torch.cuda.current_stream().synchronize()

    # ------------------------------------------------------------------ plan helpers
    @staticmethod
    def _addr(t: Optional[torch.Tensor], off_elems: int = 0) -> Optional[int]:
        if t is None:
            return None
        return t.data_ptr() + off_elems * t.element_size()

    def _buf(self, plan: Plan, name: str, shape, dtype=torch.float32, zero: bool = False) -> torch.Tensor:
        """A view of the pool entry ``name`` (grown when this shape needs more).  ``zero``: the plan relies on the
        regions no kernel writes (channel / row padding) reading as zero - see ``activate``."""
        assert name not in plan.buffers, name
        nbytes = int(math.prod(shape)) * dtype.itemsize
        t = self.pool.get(name)
        if t is None or t.numel() < nbytes:
            self.pool_grew = self.pool_grew or t is not None
            t = torch.empty((nbytes + 255) // 256 * 256, dtype=torch.uint8, device=self.dev)
            self.pool[name] = t
        v = t[:nbytes].view(dtype).view(shape)
        plan.buffers[name] = v
        if zero:
            plan.zero_bufs.append(v)
        return v

    def activate(self, plan: Plan):
        """Make ``plan`` the owner of the shared workspace: another shape may have written where this plan expects
        zero padding, so those buffers are cleared when the active plan changes (a few memsets, < 1 % of a forward)."""
        if self.active_plan is not plan:
            for t in plan.zero_bufs:
                t.zero_()
            self.active_plan = plan

    def _stats(self, plan: Plan, B: int, C: int) -> int:
        """Carve a zero-initialised (B, C, 2) double slot; returns its device address."""
        need = B * C * 2
        arenas = plan.arenas
        if not arenas or plan._arena_used + need > arenas[-1].numel():
            idx, size = len(arenas), max(need, B * 2 * 8192)
            if idx == len(self.arena_pool):
                self.arena_pool.append(torch.zeros(size, dtype=torch.float64, device=self.dev))
            elif self.arena_pool[idx].numel() < size:
                self.pool_grew = True
                self.arena_pool[idx] = torch.zeros(size, dtype=torch.float64, device=self.dev)
            arenas.append(self.arena_pool[idx][:size])
            plan._arena_used = 0
        a = arenas[-1]
        addr = a.data_ptr() + plan._arena_used * 8
        plan._arena_used += need
        return addr

    def _act_dtype(self):
        return torch.bfloat16 if self.tc else torch.float32

    def _act_pitch(self, C: int) -> int:
        if self.tc:
            return _ceil(C, 64)
        return _ceil(C, 4) if self.fp32_conv != "ffma" else C    # TMA needs 16-byte row pitches

    def _conv(self, plan: Plan, pc: PreparedConv, x_addr: int, B: int, L_in: int, x_pitch: int, M: int, *,
              dil: int = 1, pad_left: int = 0, stride: int = 1, L_true: Optional[int] = None,
              L_out: Optional[int] = None, left_pad: int = 0, up_off: int = 0, out_addr: Optional[int] = None,
              out_pitch: int = 0, use_bias: bool = True, res=None, res2=None, scale: float = 1.0,
              stats_addr: Optional[int] = None, out_act=None, force_f32: bool = False, block_n: int = 0):
        """res = (addr, pitch, shift, L_res); res2 = (addr, pitch); out_act = (addr, pitch, is_bf16, slope)."""
        d = ConvDesc()
        d.B, d.L_in, d.C_in, d.x_pitch = B, L_in, pc.C_in, x_pitch
        d.taps, d.dil, d.pad_left, d.stride = pc.taps, dil, pad_left, stride
        use_tc = self.tc and not force_f32 and pc.w16 is not None
        d.Cw = pc.Cw16 if use_tc else pc.Cw32
        d.M, d.N = M, pc.N
        d.C_out, d.up, d.up_off = pc.C_out, pc.up, up_off
        d.L_true = M if L_true is None else L_true
        d.L_out = d.L_true if L_out is None else L_out
        d.left_pad = left_pad
        d.bias = self._addr(pc.bias) if (use_bias and pc.bias is not None) else None
        if res is not None:
            d.res, d.res_pitch, d.res_shift, d.L_res = res[:4]
            d.res_is_bf16 = int(len(res) > 4 and res[4])
        if res2 is not None:
            d.res2, d.res2_pitch = res2
        d.scale = scale
        d.out, d.out_pitch = out_addr, out_pitch
        d.stats = stats_addr
        if out_act is not None:
            d.out_act, d.act_pitch, d.act_is_bf16, d.act_slope = out_act
        d.lens, d.len_mul, d.len_add = plan.lens(d.L_true)
        use_tf32 = (not self.tc) and self.fp32_conv != "ffma" and not force_f32 and stride == 1 and \
            pc.w32hi is not None and x_pitch % 4 == 0 and x_addr % 16 == 0
        if use_tf32:
            d.Cw = pc.CwT
        plan.keep.append(d)
        plan.conv_ops[len(plan.ops)] = d
        if use_tf32:
            plan.add(self.lib.kk_conv1d_tf32_tc, x_addr, self._addr(pc.w32hi), self._addr(pc.w32lo), C.byref(d),
                     3 if self.fp32_conv == "tf32x3" else 1)
        elif use_tc:
            if block_n == 0:
                block_n = self._auto_block_n(pc)
            plan.add(self.lib.kk_conv1d_bf16_tc, x_addr, self._addr(pc.w16), C.byref(d), block_n)
        else:
            assert pc.w32 is not None, pc.name
            plan.add(self.lib.kk_conv1d_f32, x_addr, self._addr(pc.w32), C.byref(d))
        return d

    def _auto_block_n(self, pc: PreparedConv) -> int:
        """N tile of the tensor-core kernel (0 = library default: 128 for N > 64).  Measured per layer at B=64 x T=320
        (experiments/ab_bn_ops.sh): a 256-wide tile is 20-25 % faster on the front convolutions (N = 1024: 210 -> 162 us),
        on decode.3's 1090 -> 512 convolutions and on noise_convs[0] (N = 256), slower on the 512 -> 512 conv2 of
        decode.3 (fewer, heavier epilogue tiles: 4.3 waves of 148 CTAs) and on the polyphase ups.1 (N = 768)."""
        if pc.N % 256 == 0 and (pc.N >= 1024 or pc.N == 256 or (pc.N == 512 and pc.C_in > 512)):
            return 256
        return 0

    def _conv_fused(self, plan: Plan, pc: PreparedConv, x: torch.Tensor, x_pitch: int, stats_addr: Optional[int],
                    ada_prefix: Optional[str], act: int, slope: float, alpha: Optional[torch.Tensor], h_all, B: int,
                    L_: int, *, dil: int = 1, pad_left: int = 0, out_addr: Optional[int] = None, out_pitch: int = 0,
                    res=None, res2=None, scale: float = 1.0, out_stats: Optional[int] = None, out_act=None,
                    out_add=None):
        """AdaIN + activation + conv in one tensor-core launch (kk_conv1d_adain_bf16_tc).  out_add = (addr, pitch):
        the result is also accumulated into that tensor (TMA reduce-add)."""
        a = AdainIn()
        a.x, a.x_is_bf16, a.x_pitch = self._addr(x), int(x.dtype == torch.bfloat16), x_pitch
        a.stats = stats_addr
        a.h = self._addr(h_all, self.ada_off[ada_prefix]) if ada_prefix is not None else None
        a.h_stride = self.ada_R
        a.alpha = self._addr(alpha)
        a.act, a.slope = act, slope
        d = ConvDesc()
        d.B, d.L_in, d.C_in, d.x_pitch = B, L_, pc.C_in, x_pitch
        d.taps, d.dil, d.pad_left, d.stride = pc.taps, dil, pad_left, 1
        d.Cw = pc.Cw16
        d.M, d.N = L_, pc.N
        d.C_out, d.up, d.up_off = pc.C_out, 1, 0
        d.L_true = d.L_out = L_
        d.left_pad = 0
        d.bias = self._addr(pc.bias) if pc.bias is not None else None
        if res is not None:
            d.res, d.res_pitch, d.res_shift, d.L_res = res[:4]
            d.res_is_bf16 = int(len(res) > 4 and res[4])
        if res2 is not None:
            d.res2, d.res2_pitch = res2
        d.scale = scale
        d.out, d.out_pitch = out_addr, out_pitch
        d.stats = out_stats
        if out_act is not None:
            d.out_act, d.act_pitch, d.act_is_bf16, d.act_slope = out_act
        if out_add is not None:
            d.out_add, d.add_pitch = out_add
        d.lens, d.len_mul, d.len_add = plan.lens(L_)
        plan.keep.extend([a, d])
        plan.conv_ops[len(plan.ops)] = d
        plan.add(self.lib.kk_conv1d_adain_bf16_tc, C.byref(a), self._addr(pc.w16), C.byref(d))
        return d

    def _ada(self, plan: Plan, x_addr: int, x_pitch: int, stats_addr: Optional[int], ada_prefix: Optional[str],
             act: int, slope: float, alpha: Optional[torch.Tensor], out: torch.Tensor, out_pitch: int, B: int, L_: int,
             C_: int, h_all: torch.Tensor, pool: Optional[str] = None):
        h_addr = self._addr(h_all, self.ada_off[ada_prefix]) if ada_prefix is not None else None
        plan.add(self.lib.kk_adain_act, x_addr, x_pitch, stats_addr, h_addr, self.ada_R,
                 self._addr(alpha), act, slope, self._addr(out), 1 if out.dtype == torch.bfloat16 else 0, out_pitch,
                 B, L_, C_, 1 if pool else 0, self._addr(self.pools[pool][0]) if pool else None,
                 self._addr(self.pools[pool][1]) if pool else None, *plan.lens(L_))

    def _tap(self, plan: Plan, name: str, t: torch.Tensor, view: Optional[Callable] = None):
        plan.taps[name] = (t, view)
        plan._push((None, (name,)))

    # ------------------------------------------------------------------ blocks
    def _adain_resblock1(self, plan: Plan, prefix: str, B: int, L_: int, C_: int, k: int, dils, h_all,
                         x_in: torch.Tensor, stats_in: int, work: torch.Tensor, act: torch.Tensor, y: torch.Tensor,
                         final: dict, x16_in: Optional[torch.Tensor] = None):
        """AdaINResBlock1 (istftnet.py:341-396).  x_in (B,L,C) fp32 with stats_in known; the running stream lives
        in ``work`` after the first iteration (work may be x_in itself: fully in place).  ``final`` gives the
        last conv2 epilogue: dict(out_addr, out_pitch, res2, scale, stats_addr, out_act).  Fused path only:
        ``out_add`` = (addr, pitch) accumulates the block's output into the MRF sum instead of reading it back as
        a second residual; ``fold_into`` = (addr, pitch) of the running MRF sum S for the LAST block of the sum:
        its stream x2 is reduce-added into S by the epilogue that produces it (S' = S + x2), and the final conv
        then takes S' as its one residual: out = (S' + conv) * scale  ==  (S + x3) * scale."""
        ap = self._act_pitch(C_)
        snake = L.KK_ACT_SNAKE_FAST if self.tc else L.KK_ACT_SNAKE  # MUFU sine only when the result is rounded to bf16
        cur, cur_stats = x_in, stats_in
        if self.fused and C_ in (128, 256):
            # tensor-core mode: AdaIN + Snake run inside the conv (A-operand transform); the conv1 output y is kept
            # in bf16 (it is only ever normalised, activated and rounded to bf16 again), the residual stream in fp32
            ybf = plan.buffers.get(f"ybf.{C_}")
            if ybf is None:
                ybf = self._buf(plan, f"ybf.{C_}", (B, L_, C_), torch.bfloat16)
            # bf16 residual stream (res16, default: the 128-channel stage): the running x is written by conv2 as bf16 only (statistics
            # still from the fp32 accumulators) and read back as bf16 by the next conv1 (input) and conv2 (residual);
            # x16_in = bf16 copy of x_in from its producer.  The block's output and the MRF sum stay fp32.
            s16 = x16_in is not None
            w16 = None
            if s16:
                w16 = plan.buffers.get(f"w16.{C_}")
                if w16 is None:
                    w16 = self._buf(plan, f"w16.{C_}", (B, L_, C_), torch.bfloat16)
                cur = x16_in
            for it in range(3):
                d = dils[it]
                c1 = self.convs[f"{prefix}.convs1.{it}"]
                c2 = self.convs[f"{prefix}.convs2.{it}"]
                st_y = self._stats(plan, B, C_)
                self._conv_fused(plan, c1, cur, C_, cur_stats, f"{prefix}.adain1.{it}", snake, 0.0,
                                 self.w[f"{prefix}.alpha1.{it}"], h_all, B, L_, dil=d, pad_left=(k * d - d) // 2,
                                 out_stats=st_y, out_act=(self._addr(ybf), C_, 1, 1.0))
                res = (self._addr(cur), C_, 0, L_, s16)
                fold = final.get("fold_into")
                if it < 2:
                    st_n = self._stats(plan, B, C_)
                    self._conv_fused(plan, c2, ybf, C_, st_y, f"{prefix}.adain2.{it}", snake, 0.0,
                                     self.w[f"{prefix}.alpha2.{it}"], h_all, B, L_, pad_left=(k - 1) // 2,
                                     out_addr=None if s16 else self._addr(work), out_pitch=C_, res=res, out_stats=st_n,
                                     out_act=(self._addr(w16), C_, 1, 1.0) if s16 else None,
                                     out_add=fold if it == 1 else None)
                    cur, cur_stats = (w16 if s16 else work), st_n
                else:
                    if fold is not None:
                        res = (fold[0], fold[1], 0, L_)
                    self._conv_fused(plan, c2, ybf, C_, st_y, f"{prefix}.adain2.{it}", snake, 0.0,
                                     self.w[f"{prefix}.alpha2.{it}"], h_all, B, L_, pad_left=(k - 1) // 2,
                                     out_addr=final.get("out_addr"), out_pitch=final.get("out_pitch", C_), res=res,
                                     res2=final.get("res2"), scale=final.get("scale", 1.0),
                                     out_stats=final.get("stats_addr"), out_act=final.get("out_act"),
                                     out_add=final.get("out_add"))
            return
        for it in range(3):
            d = dils[it]
            c1 = self.convs[f"{prefix}.convs1.{it}"]
            c2 = self.convs[f"{prefix}.convs2.{it}"]
            self._ada(plan, self._addr(cur), C_, cur_stats, f"{prefix}.adain1.{it}", snake, 0.0,
                      self.w[f"{prefix}.alpha1.{it}"], act, ap, B, L_, C_, h_all)
            st_y = self._stats(plan, B, C_)
            self._conv(plan, c1, self._addr(act), B, L_, ap, L_, dil=d, pad_left=(k * d - d) // 2,
                       out_addr=self._addr(y), out_pitch=C_, stats_addr=st_y)
            self._ada(plan, self._addr(y), C_, st_y, f"{prefix}.adain2.{it}", snake, 0.0,
                      self.w[f"{prefix}.alpha2.{it}"], act, ap, B, L_, C_, h_all)
            res = (self._addr(cur), C_, 0, L_)
            if it < 2:
                st_n = self._stats(plan, B, C_)
                self._conv(plan, c2, self._addr(act), B, L_, ap, L_, dil=1, pad_left=(k - 1) // 2,
                           out_addr=self._addr(work), out_pitch=C_, res=res, stats_addr=st_n)
                cur, cur_stats = work, st_n
            else:
                self._conv(plan, c2, self._addr(act), B, L_, ap, L_, dil=1, pad_left=(k - 1) // 2,
                           out_addr=final.get("out_addr"), out_pitch=final.get("out_pitch", C_), res=res,
                           res2=final.get("res2"), scale=final.get("scale", 1.0),
                           stats_addr=final.get("stats_addr"), out_act=final.get("out_act"))

    def _adain_resblk1d(self, plan: Plan, prefix: str, B: int, L_: int, C_in: int, C_out: int, h_all,
                        x_in_addr: int, x_pitch: int, out_addr: int, out_pitch: int, upsample: bool = False,
                        out_act=None, name: str = ""):
        """AdainResBlk1d (istftnet.py:825-899) on a channels-last fp32 input (B, L, x_pitch)."""
        Lo = 2 * L_ if upsample else L_
        adt = self._act_dtype()
        api, apo = self._act_pitch(C_in), self._act_pitch(C_out)
        st_in = self._stats(plan, B, C_in)
        plan.add(self.lib.kk_instnorm_stats, x_in_addr, st_in, B, L_, C_in, x_pitch, *plan.lens(L_))
        act1 = self._buf(plan, f"{name}.act1", (B, Lo, api), adt, zero=True)
        self._ada(plan, x_in_addr, x_pitch, st_in, f"{prefix}.norm1", L.KK_ACT_LEAKY, 0.2, None, act1, api, B, L_, C_in,
                  h_all, pool=prefix if upsample else None)
        # shortcut: [nearest x2] -> 1x1 conv without bias; the 1x1 commutes with the repeat, so it runs at L.
        # Without a learned shortcut (dim_in == dim_out, istftnet.py:840) the input itself is the residual.
        if f"{prefix}.conv1x1" in self.convs:
            xin_act = self._buf(plan, f"{name}.xin_act", (B, L_, api), adt, zero=True)
            plan.add(self.lib.kk_convert_act, x_in_addr, x_pitch, self._addr(xin_act), 1 if self.tc else 0, api,
                     B * L_, C_in, 1.0)
            sc = self._buf(plan, f"{name}.sc", (B, L_, C_out))
            self._conv(plan, self.convs[f"{prefix}.conv1x1"], self._addr(xin_act), B, L_, api, L_, use_bias=False,
                       out_addr=self._addr(sc), out_pitch=C_out)
            sc_res = (self._addr(sc), C_out, 1 if upsample else 0, L_)
        else:
            assert C_in == C_out, (prefix, C_in, C_out)
            sc_res = (x_in_addr, x_pitch, 1 if upsample else 0, L_)
        y1 = self._buf(plan, f"{name}.y1", (B, Lo, C_out))
        st1 = self._stats(plan, B, C_out)
        self._conv(plan, self.convs[f"{prefix}.conv1"], self._addr(act1), B, Lo, api, Lo, pad_left=1,
                   out_addr=self._addr(y1), out_pitch=C_out, stats_addr=st1)
        act2 = self._buf(plan, f"{name}.act2", (B, Lo, apo), adt, zero=True)
        self._ada(plan, self._addr(y1), C_out, st1, f"{prefix}.norm2", L.KK_ACT_LEAKY, 0.2, None, act2, apo, B, Lo,
                  C_out, h_all)
        self._conv(plan, self.convs[f"{prefix}.conv2"], self._addr(act2), B, Lo, apo, Lo, pad_left=1,
                   out_addr=out_addr, out_pitch=out_pitch, res=sc_res, scale=1.0 / math.sqrt(2.0), out_act=out_act)

    # ------------------------------------------------------------------ the whole decoder
    def build_plan(self, B: int, T: int, ragged: bool = False) -> Plan:
        """Decoder.__call__ (istftnet.py:947-963) + Generator.__call__ (:769-807) for a fixed (B, T).  ``ragged``: the
        batch is padded to T frames and every kernel takes the per-utterance frame counts (buffer ``in.lens``)."""
        cfg, lib = self.cfg, self.lib
        plan = Plan()
        T2, up = 2 * T, self.up_scale
        S = T2 * up
        r0, r1 = cfg["upsample_rates"]
        L20, L3 = T2 * r0, T2 * r0 * r1 + 1
        C0 = cfg["upsample_initial_channel"]
        Cs = [C0 // 2, C0 // 4]
        f32 = torch.float32
        adt = self._act_dtype()
        din = cfg["dim_in"]
        # ---- inputs (persistent so a CUDA graph can replay the plan)
        asr = self._buf(plan, "in.asr", (B, din, T))
        f0c = self._buf(plan, "in.F0_curve", (B, T2))
        nc = self._buf(plan, "in.N", (B, T2))
        s = self._buf(plan, "in.s", (B, cfg["style_dim"]))
        if ragged:
            plan.lens_addr = self._addr(self._buf(plan, "in.lens", (B,), torch.int32))
            plan.levels = {T: (1, 0), T2: (2, 0), L20: (r0 * 2, 0), L3 - 1: (r0 * r1 * 2, 0), L3: (r0 * r1 * 2, 1)}
            assert len(plan.levels) == 5, "level row counts must be distinct"
        lens_addr = plan.lens_addr
        h_all = self._buf(plan, "h_all", (B, self.ada_R))
        plan.stage("front")
        plan.add(lib.kk_style_affine, self._addr(s), self._addr(self.ada_W), self._addr(self.ada_b), self._addr(h_all),
                 B, self.ada_R, cfg["style_dim"])
        plan.fork_index = len(plan.ops)
        # ---- front: x = cat[asr, F0, N] (:951); xcat = cat[x(1024), asr_res(64), F0, N] (:957)
        c_cat0, c_cat = din + 2, 1024 + 64 + 2
        p_cat = (c_cat + 3) // 4 * 4      # 16-byte row pitch: the conv epilogues then store xcat by TMA
        p_cat0 = (c_cat0 + 3) // 4 * 4    # same for the first concatenation (vectorised elementwise kernels)
        xcat0 = self._buf(plan, "xcat0", (B, T, p_cat0), zero=True)
        xcat = self._buf(plan, "xcat", (B, T, p_cat))
        plan.add(lib.kk_front_assemble, self._addr(asr), self._addr(f0c), self._addr(nc), self._addr(self.f0_w3b),
                 self._addr(self.n_w3b), self._addr(xcat0), p_cat0, 0, din, din + 1, B, din, T, lens_addr, launches=2)
        plan.add(lib.kk_front_assemble, None, self._addr(f0c), self._addr(nc), self._addr(self.f0_w3b),
                 self._addr(self.n_w3b), self._addr(xcat), p_cat, 0, 1024 + 64, 1024 + 65, B, din, T, lens_addr)
        self._conv(plan, self.convs["asr_res.0"], self._addr(xcat0), B, T, p_cat0, T,
                   out_addr=self._addr(xcat, 1024), out_pitch=p_cat, force_f32=True)  # :953
        self._adain_resblk1d(plan, "encode", B, T, c_cat0, 1024, h_all, self._addr(xcat0), p_cat0, self._addr(xcat),
                             p_cat, name="encode")
        self._tap(plan, "encode", xcat, lambda t: t[:, :, :1024].transpose(1, 2))
        for i in range(3):
            self._adain_resblk1d(plan, f"decode.{i}", B, T, c_cat, 1024, h_all, self._addr(xcat), p_cat,
                                 self._addr(xcat), p_cat, name=f"decode{i}")
            self._tap(plan, f"decode{i}", xcat, lambda t: t[:, :, :1024].transpose(1, 2))
        x3 = self._buf(plan, "x3", (B, T2, C0))
        a_prev = self._buf(plan, "a0", (B, T2, self._act_pitch(C0)), adt, zero=True)
        self._adain_resblk1d(plan, "decode.3", B, T, c_cat, C0, h_all, self._addr(xcat), p_cat, self._addr(x3), C0,
                             upsample=True, out_act=(self._addr(a_prev), self._act_pitch(C0), int(self.tc), 0.1),
                             name="decode3")
        self._tap(plan, "decode3", x3, lambda t: t.transpose(1, 2))
        # ---- harmonic source + STFT(20,5) (:770-775)
        plan.cur_branch = 1             # ---- noise branch (needs F0 and h_all only)
        plan.stage("source")
        phase_ws = self._buf(plan, "phase_ws", (B, 9, T2))
        har_source = self._buf(plan, "har_source", (B, S))
        har = self._buf(plan, "har", (B, L3, 24))
        plan.source_op_index = len(plan.ops)
        # the Philox seed lives in device memory (written before every run / graph replay): a captured graph then
        # draws fresh noise per call like the reference's mx.random.normal (istftnet.py:620)
        seed_buf = self._buf(plan, "seed", (1,), torch.int64)
        plan.add(lib.kk_hnsf_source, self._addr(f0c), None, 0, self._addr(seed_buf), self._addr(self.lin_w), self._addr(self.lin_b),
                 self._addr(phase_ws), self._addr(har_source), B, T2, up, lens_addr, launches=2)
        har16 = None
        har_rows = r1 * (L20 + 1)       # rows of the bf16 copy: whole 6-frame rows for the stride-6 noise conv
        if self.tc:
            har16 = self._buf(plan, "har16", (B, har_rows, 24), torch.bfloat16, zero=True)
            # the convolutions read the bf16 rows only; the fp32 copy (row-strided scalar stores, 236 MB at B=64 x
            # T=320) is written when a caller asks for the intermediate taps (Decoder.__call__ patches argument 1)
            plan.stft_op_index = len(plan.ops)
            plan.har_addr = self._addr(har)
            plan.add(lib.kk_stft_n20_bf16, self._addr(har_source), self._addr(har), 24, self._addr(har16), 24, har_rows,
                     B, S, lens_addr, S // T)
        else:
            plan.add(lib.kk_stft_n20, self._addr(har_source), self._addr(har), 24, B, S, lens_addr, S // T)
        self._tap(plan, "har_source", har_source)
        self._tap(plan, "har", har, lambda t: t[:, :, :22].transpose(1, 2))
        plan.cur_branch = 0
        # ---- generator stages (:776-796)
        g = "generator"
        n_k = len(cfg["resblock_kernel_sizes"])
        Ls = [L20, L3]
        L_prev, C_prev = T2, C0
        for i in range(2):
            Li, Ci = Ls[i], Cs[i]
            plan.stage(f"stage{i}")
            plan.cur_branch = 1         # noise_convs[i] + noise_res[i]: still the noise branch
            ap = self._act_pitch(Ci)
            nconv = self.convs[f"{g}.noise_convs.{i}"]
            st_xs = self._stats(plan, B, Ci)
            fused_i = self.fused and Ci in (128, 256)
            s16 = fused_i and (self.res16 >= 2 or (self.res16 == 1 and Ci == 128))
            # bf16 stream: the noise branch x_source exists only as bf16 - written by noise_convs, rewritten by the last
            # conv of noise_res (its input copy is free by then) and read back as the bf16 residual of ups (x + x_source,
            # istftnet.py:788): 0.63 GB instead of 1.26 GB written and read at B=64 x T=320
            xs = None if s16 else self._buf(plan, f"g{i}.xs", (B, Li, Ci))
            xs16 = self._buf(plan, f"g{i}.xs16", (B, Li, Ci), torch.bfloat16) if s16 else None
            x16 = self._buf(plan, f"g{i}.x16", (B, Li, Ci), torch.bfloat16) if s16 else None
            xs_out = None if s16 else self._addr(xs)     # bf16 stream: the noise conv's result is only ever read as bf16
            if self.tc:
                ntc = self.convs[f"{g}.noise_convs.{i}.tc"]
                oa = (self._addr(xs16), Ci, 1, 1.0) if s16 else None
                if i == 0:   # (B, L20+1 rows of r1 frames, r1*24 channels), 3 taps, left pad 1
                    self._conv(plan, ntc, self._addr(har16), B, L20 + 1, r1 * 24, Li, pad_left=1,
                               out_addr=xs_out, out_pitch=Ci, stats_addr=st_xs, out_act=oa)
                else:        # 1x1 over the frames; rows >= L3 of har16 are never read (M = L3)
                    self._conv(plan, ntc, self._addr(har16), B, har_rows, 24, Li, out_addr=xs_out,
                               out_pitch=Ci, stats_addr=st_xs, out_act=oa)
            elif i == 0:
                stride_f0 = r1
                self._conv(plan, nconv, self._addr(har), B, L3, 24, Li, stride=stride_f0, pad_left=(stride_f0 + 1) // 2,
                           out_addr=self._addr(xs), out_pitch=Ci, stats_addr=st_xs, force_f32=True)
            else:
                self._conv(plan, nconv, self._addr(har), B, L3, 24, Li, out_addr=self._addr(xs), out_pitch=Ci,
                           stats_addr=st_xs, force_f32=True)
            self._tap(plan, f"noise_conv{i}", xs16 if s16 else xs, lambda t: t.transpose(1, 2))
            act = None if fused_i else self._buf(plan, f"g{i}.act", (B, Li, ap), adt, zero=True)
            y = None if fused_i else self._buf(plan, f"g{i}.y", (B, Li, Ci))
            self._adain_resblock1(plan, f"{g}.noise_res.{i}", B, Li, Ci, 7 if i == 0 else 11, (1, 3, 5), h_all, xs,
                                  st_xs, xs, act, y,
                                  dict(out_act=(self._addr(xs16), Ci, 1, 1.0)) if s16 else dict(out_addr=self._addr(xs), out_pitch=Ci),
                                  x16_in=xs16)
            self._tap(plan, f"noise_res{i}", xs16 if s16 else xs, lambda t: t.transpose(1, 2))
            plan._push((None, ("@done", i)))
            plan.cur_branch = 0
            plan._push((None, ("@join", i)))
            # ups[i]: polyphase transposed conv, + x_source fused as residual (:783-788)
            u = cfg["upsample_rates"][i]
            kk_ = cfg["upsample_kernel_sizes"][i]
            x = None if s16 else self._buf(plan, f"g{i}.x", (B, Li, Ci))
            st_x = self._stats(plan, B, Ci)
            last = i == 1
            self._conv(plan, self.convs[f"{g}.ups.{i}"], self._addr(a_prev), B, L_prev, self._act_pitch(C_prev),
                       L_prev + 1, pad_left=1, up_off=(kk_ - u) // 2, L_true=L_prev * u, L_out=Li,
                       left_pad=1 if last else 0, out_addr=None if s16 else self._addr(x),
                       out_pitch=Ci, res=(self._addr(xs16), Ci, 0, Li, True) if s16 else (self._addr(xs), Ci, 0, Li),
                       stats_addr=st_x,
                       out_act=(self._addr(x16), Ci, 1, 1.0) if s16 else None)
            self._tap(plan, f"x_stage{i}", x16 if s16 else x, lambda t: t.transpose(1, 2))
            # MRF: mean of three AdaINResBlock1 on the same input (:790-796)
            work = None if s16 else self._buf(plan, f"g{i}.work", (B, Li, Ci))
            acc = self._buf(plan, f"g{i}.mrf", (B, Li, Ci))
            a_next = self._buf(plan, f"a{i + 1}", (B, Li, ap), adt, zero=True)
            for j in range(n_k):
                final = dict(out_addr=self._addr(acc), out_pitch=Ci)
                if j > 0 and fused_i:       # the sum over resblocks is carried by TMA reduce-adds
                    if j < n_k - 1:
                        final = dict(out_add=(self._addr(acc), Ci))
                    else:
                        final["fold_into"] = (self._addr(acc), Ci)
                elif j > 0:
                    final["res2"] = (self._addr(acc), Ci)
                if j == n_k - 1:
                    final["scale"] = 1.0 / n_k
                    final["out_act"] = (self._addr(a_next), ap, int(self.tc), 0.1 if i == 0 else 0.01)
                self._adain_resblock1(plan, f"{g}.resblocks.{i * n_k + j}", B, Li, Ci, cfg["resblock_kernel_sizes"][j],
                                      cfg["resblock_dilation_sizes"][j], h_all, x, st_x, work, act, y, final,
                                      x16_in=x16)
            self._tap(plan, f"mrf{i}", acc, lambda t: t.transpose(1, 2))
            a_prev, L_prev, C_prev = a_next, Li, Ci
        # ---- conv_post + iSTFT head (:798-806)
        plan.stage("post")
        ypost = self._buf(plan, "ypost", (B, L3, 24))
        pc_post = self.convs[f"{g}.conv_post"]
        if self.fused and Cs[1] == 128 and pc_post.N <= 32 and self._act_pitch(Cs[1]) == 128:
            # row-shifted-descriptor kernel: the input tile + halo is read once instead of once per tap
            self._conv_fused(plan, pc_post, a_prev, 128, None, None, L.KK_ACT_NONE, 0.0, None, h_all, B, L3, pad_left=3,
                             out_addr=self._addr(ypost), out_pitch=24)
        else:
            self._conv(plan, self.convs[f"{g}.conv_post"], self._addr(a_prev), B, L3, self._act_pitch(Cs[1]), L3, pad_left=3,
                       out_addr=self._addr(ypost), out_pitch=24)
        self._tap(plan, "conv_post", ypost, lambda t: t[:, :, :22].transpose(1, 2))
        audio = self._buf(plan, "audio", (B, 1, S))
        plan.add(lib.kk_istft_head_n20, self._addr(ypost), 24, self._addr(audio), B, L3, *plan.lens(L3))
        plan.shape = (B, T)
        return plan
