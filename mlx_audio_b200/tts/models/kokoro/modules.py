"""B200-native drop-in for the F0 / energy heads of mlx_audio/tts/models/kokoro/modules.py ``ProsodyPredictor``
(SURVEY 8(f) row 2): ``F0Ntrain(x, s)`` (modules.py:355-377) - the producer of the decoder's ``F0_curve`` / ``N`` inputs.

    x (B, d_hid + style_dim, L) --shared bi-LSTM--> (B, L, d_hid)
        --3 x AdainResBlk1d (d_hid -> d_hid, d_hid -> d_hid/2 with x2 upsampling, d_hid/2 -> d_hid/2)--> (B, d_hid/2, 2L)
        --1x1 Conv1d--> F0 (B, 2L);   the same again with the ``N`` blocks -> N (B, 2L)

The six AdainResBlk1d (istftnet.py:825-899, built at modules.py:299-337 - blocks 0 and 2 have no learned shortcut) and
the two 1x1 projections run on the same hand-written kernels as the decoder front (launch plan: ``ProsodyEngine``,
reusing ``DecoderEngine._adain_resblk1d``).  The recurrent ``shared`` LSTM (modules.py:93-286, gate order i, f, g, o)
is native as well (round 2; it was a cuDNN module): one tensor-core GEMM for the input projections of all steps, then
``kk_lstm_bidir_f32`` (``csrc/lstm.cu``) - clusters of 8 CTAs that keep the recurrent weights in shared memory and
exchange h once per step.  It may still be skipped by passing its output as ``hidden=``.  The rest of ``ProsodyPredictor`` (DurationEncoder, duration_proj)
is outside the rebuilt path.  Parameter names are the reference's (``F0.0.conv1.weight_v``, ``N_proj.weight`` ...),
with or without the ``predictor.`` prefix.  No CPU path.
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Tuple

import numpy as np
import torch

from .... import _lib as L
from ....synth import prosody_param_specs
from ..base import check_array_shape
from .engine import DecoderEngine, Plan, PreparedConv


class ProsodyEngine(DecoderEngine):
    """Launch plans of the convolutional part of ``F0Ntrain`` for fixed (B, L)."""

    def _check_cfg(self):
        self.d_hid = int(self.cfg["d_hid"])

    def _prepare(self):
        w = self.w
        self._prepare_adain()
        for head in ("F0", "N"):
            for i in range(3):
                self._prep_resblk1d(f"{head}.{i}")
            # nn.Conv1d(d_hid // 2, 1, 1) (modules.py:339-340): plain weight (1, 1, C) + bias, fp32 CUDA-core conv
            self.convs[f"{head}_proj"] = PreparedConv(f"{head}_proj", w[f"{head}_proj.weight"].contiguous(),
                                                      w[f"{head}_proj.bias"], need_bf16=False, need_f32=True)
        torch.cuda.current_stream().synchronize()

    def build_plan(self, B: int, L_: int, ragged: bool = False) -> Plan:
        """``ragged``: the batch is padded to L_ frames and every kernel takes the per-sequence frame counts (buffer
        ``pp.in.lens``), like the decoder's plans (engine.build_plan)."""
        plan = Plan()
        d, h2 = self.d_hid, self.d_hid // 2
        if ragged:
            plan.lens_addr = self._addr(self._buf(plan, "pp.in.lens", (B,), torch.int32))
            plan.levels = {L_: (1, 0), 2 * L_: (2, 0)}
        x = self._buf(plan, "pp.in.x", (B, L_, d))            # shared-LSTM output, channels-last (modules.py:358)
        s = self._buf(plan, "pp.in.s", (B, self.cfg["style_dim"]))
        h_all = self._buf(plan, "pp.h_all", (B, self.ada_R))
        plan.add(self.lib.kk_style_affine, self._addr(s), self._addr(self.ada_W), self._addr(self.ada_b),
                 self._addr(h_all), B, self.ada_R, self.cfg["style_dim"])
        # the two heads share nothing but their inputs: N runs as a parallel branch (Plan.run_parallel, like the
        # decoder's noise branch), forked here and joined after F0's last launch
        plan.fork_index = len(plan.ops)
        for head in ("F0", "N"):
            plan.cur_branch = 1 if head == "N" else 0
            y0 = self._buf(plan, f"pp.{head}.y0", (B, L_, d))
            y1 = self._buf(plan, f"pp.{head}.y1", (B, 2 * L_, h2))
            y2 = self._buf(plan, f"pp.{head}.y2", (B, 2 * L_, h2))
            out = self._buf(plan, f"pp.out.{head}", (B, 2 * L_))
            self._adain_resblk1d(plan, f"{head}.0", B, L_, d, d, h_all, self._addr(x), d, self._addr(y0), d,
                                 name=f"pp.{head}0")
            self._adain_resblk1d(plan, f"{head}.1", B, L_, d, h2, h_all, self._addr(y0), d, self._addr(y1), h2,
                                 upsample=True, name=f"pp.{head}1")
            self._adain_resblk1d(plan, f"{head}.2", B, 2 * L_, h2, h2, h_all, self._addr(y1), h2, self._addr(y2), h2,
                                 name=f"pp.{head}2")
            self._tap(plan, f"{head}.blocks", y2, lambda t: t.transpose(1, 2))
            self._conv(plan, self.convs[f"{head}_proj"], self._addr(y2), B, 2 * L_, h2, 2 * L_,
                       out_addr=self._addr(out), out_pitch=1, force_f32=True)
        plan._push((None, ("@done", 0)))
        plan.cur_branch = 0
        plan._push((None, ("@join", 0)))
        plan.shape = (B, L_)
        return plan


def _to_cuda_f32(x, device) -> torch.Tensor:
    if isinstance(x, torch.Tensor):
        return x.to(device=device, dtype=torch.float32, non_blocking=True)
    return torch.as_tensor(np.asarray(x), dtype=torch.float32).to(device, non_blocking=True)


class ProsodyPredictor:
    """``ProsodyPredictor(style_dim, d_hid, nlayers, max_dur, dropout)`` (modules.py:288-341) restricted to what
    ``F0Ntrain`` touches: ``shared``, ``F0``, ``N``, ``F0_proj``, ``N_proj``."""

    def __init__(self, style_dim: int = 128, d_hid: int = 512, nlayers: int = 3, max_dur: int = 50,
                 dropout: float = 0.1, dtype: str = "bf16", device: Optional[str] = None):
        if dtype in ("float32", torch.float32):
            dtype = "fp32"
        if dtype in ("bfloat16", torch.bfloat16):
            dtype = "bf16"
        self.cfg = dict(style_dim=style_dim, d_hid=d_hid)
        self.mode = dtype
        self.device = torch.device(device or "cuda")
        self._specs = {n[len("predictor."):]: shp for n, shp, _ in prosody_param_specs(style_dim, d_hid)}
        self._weights: Dict[str, torch.Tensor] = {}
        self._engine: Optional[ProsodyEngine] = None
        self._lstm = None      # (input-projection GEMM, transposed recurrent weights) once the weights are loaded
        self.parallel_heads = True          # F0 and N heads on two streams (False: one stream, same launches)
        self._side: Optional[torch.cuda.Stream] = None

    # ------------------------------------------------------------------ weights
    def parameter_shapes(self) -> Dict[str, tuple]:
        return dict(self._specs)

    _LSTM_MAP = {"weight_ih_l0_reverse": "Wx_backward", "weight_hh_l0_reverse": "Wh_backward",
                 "bias_ih_l0_reverse": "bias_ih_backward", "bias_hh_l0_reverse": "bias_hh_backward",
                 "weight_ih_l0": "Wx_forward", "weight_hh_l0": "Wh_forward", "bias_ih_l0": "bias_ih_forward",
                 "bias_hh_l0": "bias_hh_forward"}                                    # kokoro.py:29-38

    @classmethod
    def sanitize(cls, key: str, weights) -> Dict[str, object]:
        """The ``predictor`` branch of Model.sanitize (kokoro.py:213-247), one key at a time -> {new_key: array}:
        ``F0_proj.weight`` / ``N_proj.weight`` are always transposed (O, I, K) -> (O, K, I) (:214-218), ``weight_v``
        unless ``check_array_shape`` says it already is (:220-224), torch LSTM names become Wx/Wh/bias_* (:24-44)."""
        def tr(a):
            return a.transpose(0, 2, 1) if isinstance(a, np.ndarray) else a.permute(0, 2, 1)
        if "F0_proj.weight" in key or "N_proj.weight" in key:
            return {key: tr(weights)}
        if "weight_v" in key:
            return {key: weights if check_array_shape(weights) else tr(weights)}
        base, _, suffix = key.rpartition(".")
        if suffix in cls._LSTM_MAP:
            return {f"{base}.{cls._LSTM_MAP[suffix]}": weights}
        return {key: weights}

    def load_weights(self, weights, strict: bool = True):
        L.require_cuda()
        items = weights.items() if isinstance(weights, dict) else weights
        got = {}
        for k, v in items:
            if k.startswith("predictor."):
                k = k[len("predictor."):]
            if k not in self._specs:
                if strict:
                    raise ValueError(f"Unexpected predictor parameter {k!r}")
                continue
            t = _to_cuda_f32(v, self.device).contiguous()
            if tuple(t.shape) != tuple(self._specs[k]):
                raise ValueError(f"Shape mismatch for {k}: expected {self._specs[k]}, got {tuple(t.shape)}")
            got[k] = t
        missing = {k for k in set(self._specs) - set(got) if not k.startswith("shared.")}
        if strict and missing:
            raise ValueError(f"Missing predictor parameters: {sorted(missing)[:5]} ... ({len(missing)} total)")
        self._weights.update(got)
        self._engine = None
        self._lstm = None
        return self

    @property
    def engine(self) -> ProsodyEngine:
        if self._engine is None:
            conv_w = {k: v for k, v in self._weights.items() if not k.startswith("shared.")}
            need = {k for k in self._specs if not k.startswith("shared.")}
            if set(conv_w) != need:
                raise L.KokoroB200Error("ProsodyPredictor weights are not loaded (call load_weights first)")
            self._engine = ProsodyEngine(conv_w, self.cfg, self.mode)
        return self._engine

    def shared(self, x_blc: torch.Tensor, lengths=None) -> torch.Tensor:
        """``self.shared`` bi-LSTM (modules.py:93-286, gate order i, f, g, o): (B, L, d_hid + style) -> (B, L, d_hid).
        Input projection of every step and both directions = one 3xTF32 tensor-core GEMM (``kk_conv1d_tf32_tc`` as a
        1x1 convolution, bias = b_ih + b_hh); the recurrence = ``kk_lstm_bidir_f32`` (thread-block clusters holding the
        recurrent weights in shared memory).  fp32 in both modes - it is ~1 ms beside a 35 ms decode.
        ``lengths`` (extension): valid steps per sequence of a zero-padded batch; the backward direction starts at each
        sequence's own end, rows past it are zeros."""
        from ...._gemm import LinearTF32x3
        lib = L.load()
        w = self._weights
        if self._lstm is None:
            if "shared.Wx_forward" not in w:
                raise L.KokoroB200Error("the `shared` LSTM weights are not loaded: pass hidden= (its output) instead")
            dev = self.device
            f32 = lambda t: t.to(device=dev, dtype=torch.float32)
            wx = torch.cat([f32(w["shared.Wx_forward"]), f32(w["shared.Wx_backward"])], 0)          # (2 * 4H, D)
            bias = torch.cat([f32(w["shared.bias_ih_forward"]) + f32(w["shared.bias_hh_forward"]),
                              f32(w["shared.bias_ih_backward"]) + f32(w["shared.bias_hh_backward"])]).contiguous()
            pc = LinearTF32x3(wx.contiguous(), bias)
            whT = torch.stack([f32(w["shared.Wh_forward"]).t(), f32(w["shared.Wh_backward"]).t()]).contiguous()
            self._lstm = (pc, whT)
        pc, whT = self._lstm
        x = x_blc.to(device=self.device, dtype=torch.float32).contiguous()
        B, L_, D = x.shape
        H = self.cfg["d_hid"] // 2
        if D != pc.n_in:
            raise ValueError(f"shared LSTM input must be (B, L, {pc.n_in}), got {tuple(x.shape)}")
        lens = None
        if lengths is not None:
            lens = torch.as_tensor(lengths).to(device=self.device, dtype=torch.int32).contiguous()
            if lens.numel() != B:
                raise ValueError(f"lengths: expected {B} entries, got {lens.numel()}")
        xw = pc(x)                                              # (B, L, 2 * 4H)
        st = L.stream()
        out = torch.empty(B, L_, 2 * H, device=self.device)
        L.check(lib.kk_lstm_bidir_f32(L.ptr(xw), L.ptr(whT), L.ptr(out), B, L_, H,
                                      None if lens is None else L.ptr(lens), st), "kk_lstm_bidir_f32")
        return out

    # ------------------------------------------------------------------ forward
    def F0Ntrain(self, x=None, s=None, *, hidden=None, capture_taps: bool = False,
                 borrow: bool = False, lengths=None) -> Tuple[torch.Tensor, torch.Tensor]:
        """modules.py:355-377.  ``x``: (B, d_hid + style_dim, L) (``en`` of kokoro.py:157-159); ``s``: (B, style_dim).
        Extension: ``hidden`` = (B, L, d_hid) output of the shared LSTM, skipping it.  Returns (F0, N), each (B, 2L):
        fresh tensors like the reference; ``borrow=True`` returns the persistent workspace views (overwritten by the
        next call).  ``lengths`` (extension): frames per sequence of a zero-padded RAGGED batch - every sequence comes
        out as if it were run alone (LSTM directions, instance-norm statistics, conv padding), zeros after 2 * length."""
        L.require_cuda()
        dev = self.device
        s = _to_cuda_f32(s, dev)
        lens = None
        if lengths is not None:
            if isinstance(lengths, torch.Tensor) and lengths.is_cuda:
                lens = lengths.to(torch.int32).contiguous()
            else:
                host = np.asarray(lengths.cpu() if isinstance(lengths, torch.Tensor) else lengths).astype(np.int64).reshape(-1)
                Lmax = (x if hidden is None else hidden).shape[2 if hidden is None else 1]
                if host.size and (host.min() < 1 or host.max() > Lmax):
                    raise ValueError(f"lengths must lie in [1, {Lmax}], got [{host.min()}, {host.max()}]")
                lens = torch.from_numpy(host.astype(np.int32)).to(dev)
        if hidden is None:
            x = _to_cuda_f32(x, dev)
            if x.dim() != 3 or x.shape[1] != self.cfg["d_hid"] + self.cfg["style_dim"]:
                raise ValueError(f"x must be (B, {self.cfg['d_hid'] + self.cfg['style_dim']}, L), got {tuple(x.shape)}")
            hidden = self.shared(x.transpose(1, 2).contiguous(), lengths=lens)
        hidden = _to_cuda_f32(hidden, dev)
        if hidden.dim() != 3 or hidden.shape[2] != self.cfg["d_hid"]:
            raise ValueError(f"hidden must be (B, L, {self.cfg['d_hid']}), got {tuple(hidden.shape)}")
        B, L_, _ = hidden.shape
        if tuple(s.shape) != (B, self.cfg["style_dim"]):
            raise ValueError(f"s must be (B, {self.cfg['style_dim']}), got {tuple(s.shape)}")
        if lens is not None and lens.numel() != B:
            raise ValueError(f"lengths: expected {B} entries, got {lens.numel()}")
        eng = self.engine
        key = (B, L_) if lens is None else (B, L_, True)
        plan = eng.plans.get(key)
        if plan is None:
            eng.pool_grew = False
            plan = eng.build_plan(B, L_, ragged=lens is not None)
            if eng.pool_grew:
                eng.plans.clear()
            eng.plans[key] = plan
        eng.activate(plan)
        if lens is not None:
            plan.buffers["pp.in.lens"].copy_(lens, non_blocking=True)
        plan.buffers["pp.in.x"].copy_(hidden, non_blocking=True)
        plan.buffers["pp.in.s"].copy_(s, non_blocking=True)
        plan.capture = capture_taps
        if self.parallel_heads and not capture_taps:
            if self._side is None:
                self._side = torch.cuda.Stream(dev)
            plan.run_parallel(self._side)
        else:
            plan.run(L.stream())
        f0, n = plan.buffers["pp.out.F0"], plan.buffers["pp.out.N"]
        return (f0, n) if borrow else (f0.clone(), n.clone())

    def launches_per_forward(self, B: int, L_: int) -> int:
        return self.engine.plans[(B, L_)].n_launches


def build_prosody_predictor(weights=None, style_dim: int = 128, d_hid: int = 512, dtype: str = "bf16",
                            device: Optional[str] = None) -> ProsodyPredictor:
    p = ProsodyPredictor(style_dim, d_hid, dtype=dtype, device=device)
    if weights is not None:
        p.load_weights(weights)
    return p
