"""B200-native drop-in for mlx_audio/tts/models/kokoro/istftnet.py ``Decoder``.

Same construction arguments, same parameter names (``weight_g`` / ``weight_v`` / ``bias`` / ``fc.weight`` /
``alpha1.N`` ...), same ``sanitize`` layout rule and the same call contract

    Decoder.__call__(asr (B,512,T), F0_curve (B,2T), N (B,2T), s (B,128)) -> (B, 1, 600*T)

(reference: istftnet.py:902-979).  Host code is Python + torch tensor plumbing only; every arithmetic step is
a hand-written sm_100a kernel reached through the C ABI of libkokoro_b200 (see engine.py for the launch
plan).  There is no CPU path: without a CUDA device or the built shared library the class raises.

Extensions over the reference signature (keyword-only, defaults keep reference behaviour):
  * ``noise``: (B, 600T, 9) standard-normal tensor replacing the reference's ``mx.random.normal`` draw in
    SineGen (istftnet.py:620) so results are reproducible / comparable with the oracle; default: on-device
    Philox RNG seeded by ``seed`` (a fresh seed per call, also under CUDA graphs: the seed is read from device memory).
  * ``borrow``: return the persistent output buffer instead of a copy (see ``__call__``).
"""
from __future__ import annotations

from typing import Dict, Optional

import numpy as np
import torch

from .... import _lib as L
from ....synth import KOKORO_DECODER, decoder_param_specs
from ..base import check_array_shape
from .engine import DecoderEngine, Plan


def get_padding(kernel_size: int, dilation: int = 1) -> int:
    """istftnet.py:14-15."""
    return int((kernel_size * dilation - dilation) / 2)


def _to_cuda_f32(x, device) -> torch.Tensor:
    if isinstance(x, torch.Tensor):
        return x.to(device=device, dtype=torch.float32, non_blocking=True)
    return torch.as_tensor(np.asarray(x), dtype=torch.float32).to(device, non_blocking=True)


class Decoder:
    """Kokoro ISTFTNet decoder (AdainResBlk1d stack + Generator + iSTFT head) on B200."""

    def __init__(self, dim_in, style_dim, dim_out, resblock_kernel_sizes, upsample_rates, upsample_initial_channel,
                 resblock_dilation_sizes, upsample_kernel_sizes, gen_istft_n_fft, gen_istft_hop_size,
                 dtype: str = "bf16", device: Optional[str] = None, engine_options: Optional[dict] = None):
        self.cfg = dict(dim_in=dim_in, style_dim=style_dim, dim_out=dim_out,
                        resblock_kernel_sizes=list(resblock_kernel_sizes), upsample_rates=list(upsample_rates),
                        upsample_initial_channel=upsample_initial_channel,
                        resblock_dilation_sizes=[list(d) for d in resblock_dilation_sizes],
                        upsample_kernel_sizes=list(upsample_kernel_sizes), gen_istft_n_fft=gen_istft_n_fft,
                        gen_istft_hop_size=gen_istft_hop_size)
        if dtype in ("float32", torch.float32):
            dtype = "fp32"
        if dtype in ("bfloat16", torch.bfloat16):
            dtype = "bf16"
        self.mode = dtype
        self.device = torch.device(device or "cuda")
        self._specs = {n[len("decoder."):]: shp for n, shp, _ in decoder_param_specs(self.cfg)}
        self._weights: Dict[str, torch.Tensor] = {}
        self._engine: Optional[DecoderEngine] = None
        self._graphs: Dict[tuple, torch.cuda.CUDAGraph] = {}
        self._engine_options = dict(engine_options or {})   # DecoderEngine arms: fused=, res16= (see engine.py)
        self.use_cuda_graph = False
        # run the generator's noise branch on a side stream (a parallel branch inside a CUDA graph).  Measured on B200
        # (bf16, graph replay): B=1 x T=200 2.10 -> 1.72 ms, B=8 3.85 -> 3.50, B=16 x T=320 9.58 -> 9.32, B=64 36.4 -> 36.3:
        # on by default; "auto" = only while B * T <= 2048 frames; False = one stream
        self.parallel_branches = True
        self._side: Optional[torch.cuda.Stream] = None
        self.seed = 0

    # ------------------------------------------------------------------ weights
    def parameter_shapes(self) -> Dict[str, tuple]:
        return dict(self._specs)

    def sanitize(self, key, weights):
        """istftnet.py:965-979: torch-layout (O, I, K) conv weights -> MLX layout (O, K, I)."""
        if "noise_convs" in key and key.endswith(".weight"):
            return weights.transpose(0, 2, 1) if isinstance(weights, np.ndarray) else weights.permute(0, 2, 1)
        if "weight_v" in key:
            if check_array_shape(weights):
                return weights
            return weights.transpose(0, 2, 1) if isinstance(weights, np.ndarray) else weights.permute(0, 2, 1)
        return weights

    def load_weights(self, weights, strict: bool = True):
        """``weights``: dict or list of (name, array); names with or without the ``decoder.`` prefix, MLX layout."""
        L.require_cuda()
        items = weights.items() if isinstance(weights, dict) else weights
        got = {}
        for k, v in items:
            if k.startswith("decoder."):
                k = k[len("decoder."):]
            if k not in self._specs:
                if strict:
                    raise ValueError(f"Unexpected decoder parameter {k!r}")
                continue
            t = _to_cuda_f32(v, self.device).contiguous()
            if tuple(t.shape) != tuple(self._specs[k]):
                raise ValueError(f"Shape mismatch for {k}: expected {self._specs[k]}, got {tuple(t.shape)}")
            got[k] = t
        missing = set(self._specs) - set(got)
        if strict and missing:
            raise ValueError(f"Missing decoder parameters: {sorted(missing)[:5]} ... ({len(missing)} total)")
        self._weights.update(got)
        self._engine = None
        self._graphs.clear()
        return self

    @property
    def engine(self) -> DecoderEngine:
        if self._engine is None:
            if len(self._weights) != len(self._specs):
                raise L.KokoroB200Error("Decoder weights are not loaded (call load_weights first)")
            self._engine = DecoderEngine(self._weights, self.cfg, self.mode, **self._engine_options)
        return self._engine

    # ------------------------------------------------------------------ forward
    def _plan(self, B: int, T: int, ragged: bool = False) -> Plan:
        eng = self.engine
        key = (B, T, ragged)
        plan = eng.plans.get(key)
        if plan is None:
            eng.pool_grew = False
            plan = eng.build_plan(B, T, ragged)
            if eng.pool_grew:
                # a workspace buffer was re-allocated: plans (and graphs) built against the old one would keep it
                # alive, so they are dropped and rebuilt on demand (call reserve(B_max, T_max) first to avoid this)
                eng.plans.clear()
                self._graphs.clear()
            eng.plans[key] = plan
        return plan

    def reserve(self, B: int, T: int, ragged: bool = False) -> None:
        """Size the shared workspace for the largest (B, T) the caller will decode (ragged / bucketed workloads)."""
        self._plan(B, T, ragged)

    def __call__(self, asr, F0_curve, N, s, *, noise=None, seed: Optional[int] = None, capture_taps: bool = False,
                 borrow: bool = False, lengths=None):
        """Returns a fresh (B, 1, 600 T) tensor like the reference (istftnet.py:963).  ``borrow=True`` hands out the
        decoder's persistent output buffer instead (zero copy): it is overwritten by the next call of ANY shape, so only
        callers that consume it before calling again (StreamingDecoder, decode_sharded, bench.py) should ask for it.

        ``lengths`` (B,) ints: a RAGGED batch - utterance b has ``lengths[b] <= T`` asr frames, the inputs are padded to
        T (the padding's content is irrelevant).  Every kernel of the one launch plan masks per utterance (padding rows
        are the convolutions' zero padding and stay out of the instance-norm statistics), so ``out[b, 0, :600 *
        lengths[b]]`` equals decoding utterance b alone - the reference decodes one utterance per call
        (kokoro.py:135-165) - and ``out[b, 0, 600 * lengths[b]:]`` is zero."""
        L.require_cuda()
        dev = self.device
        asr = _to_cuda_f32(asr, dev)
        F0_curve = _to_cuda_f32(F0_curve, dev)
        N = _to_cuda_f32(N, dev)
        s = _to_cuda_f32(s, dev)
        if asr.dim() != 3 or asr.shape[1] != self.cfg["dim_in"]:
            raise ValueError(f"asr must be (B, {self.cfg['dim_in']}, T), got {tuple(asr.shape)}")
        B, _, T = asr.shape
        if tuple(F0_curve.shape) != (B, 2 * T) or tuple(N.shape) != (B, 2 * T):
            raise ValueError(f"F0_curve and N must be (B, 2T) = ({B}, {2 * T}); got {tuple(F0_curve.shape)}, {tuple(N.shape)}")
        if tuple(s.shape) != (B, self.cfg["style_dim"]):
            raise ValueError(f"s must be (B, {self.cfg['style_dim']}), got {tuple(s.shape)}")
        ragged = lengths is not None
        if ragged:
            lens_t = torch.as_tensor(lengths).to(torch.int32).reshape(-1)
            if lens_t.numel() != B:
                raise ValueError(f"lengths must have {B} entries, got {lens_t.numel()}")
            if not lens_t.is_cuda:
                # host-resident lengths are validated; device-resident ones are trusted (reading them back would stall
                # the launch pipeline on every call) - the kernels clamp them to [0, T] (common.cuh: ragged_len)
                lo, hi = int(lens_t.min()), int(lens_t.max())
                if lo < 1 or hi > T:
                    raise ValueError(f"lengths must lie in [1, T = {T}], got [{lo}, {hi}]")
        plan = self._plan(B, T, ragged)
        self.engine.activate(plan)
        bufs = plan.buffers
        if ragged:
            bufs["in.lens"].copy_(lens_t, non_blocking=True)
        bufs["in.asr"].copy_(asr, non_blocking=True)
        bufs["in.F0_curve"].copy_(F0_curve, non_blocking=True)
        bufs["in.N"].copy_(N, non_blocking=True)
        bufs["in.s"].copy_(s, non_blocking=True)
        noise_t = None
        if noise is not None:
            noise_t = _to_cuda_f32(noise, dev).contiguous()
            S = 2 * T * self.engine.up_scale
            if tuple(noise_t.shape) != (B, S, 9):
                raise ValueError(f"noise must be (B, {S}, 9), got {tuple(noise_t.shape)}")
        self._set_source(plan, noise_t, self.seed if seed is None else seed)
        plan.capture = capture_taps
        i = getattr(plan, "stft_op_index", None)
        if i is not None:               # fp32 copy of the harmonic-source STFT: only for the taps (see engine.build_plan)
            fn, args = plan.ops[i]
            plan.ops[i] = (fn, (args[0], plan.har_addr if capture_taps else None) + tuple(args[2:]))
        par = self.parallel_branches
        par = (B * T <= 2048) if par == "auto" else bool(par)
        par = par and not capture_taps and plan.fork_index is not None
        use_graph = self.use_cuda_graph and noise_t is None and not capture_taps
        if use_graph:
            self._run_graph(plan, (B, T, ragged, par), par)
        else:
            self._run_plan(plan, par)
        if seed is None:
            self.seed += 1
        return bufs["audio"] if borrow else bufs["audio"].clone()

    def _set_source(self, plan: Plan, noise_t: Optional[torch.Tensor], seed: int):
        i = plan.source_op_index
        fn, args = plan.ops[i]
        args = list(args)
        args[1] = noise_t.data_ptr() if noise_t is not None else None
        plan.ops[i] = (fn, tuple(args))
        plan.buffers["seed"].fill_(int(seed) & 0x7FFFFFFFFFFFFFFF)   # outside any captured graph
        plan.keep_noise = noise_t

    def _run_plan(self, plan: Plan, par: bool):
        if par:
            if self._side is None:
                self._side = torch.cuda.Stream(self.device)
            plan.run_parallel(self._side)
        else:
            plan.run(L.stream())

    def _run_graph(self, plan: Plan, key, par: bool = False):
        g = self._graphs.get(key)
        if g is None:
            plan.run(L.stream())  # warm-up: sets kernel attributes outside capture
            torch.cuda.current_stream().synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._run_plan(plan, par)
            self._graphs[key] = g
        g.replay()

    def taps(self, B: int, T: int, ragged: bool = False) -> Dict[str, torch.Tensor]:
        return self._plan(B, T, ragged).captured

    def launches_per_forward(self, B: int, T: int, ragged: bool = False) -> int:
        return self._plan(B, T, ragged).n_launches


def build_decoder(weights=None, cfg: Optional[dict] = None, dtype: str = "bf16", device: Optional[str] = None,
                  engine_options: Optional[dict] = None) -> Decoder:
    """Convenience constructor: Decoder(**cfg) + load_weights."""
    dec = Decoder(**(cfg or KOKORO_DECODER), dtype=dtype, device=device, engine_options=engine_options)
    if weights is not None:
        dec.load_weights(weights)
    return dec
