#!/usr/bin/env python
"""Benchmark of the Kokoro ISTFTNet decoder hot path (BASELINE.json metric: audio-seconds generated per
wall-second at 24 kHz, decoder + iSTFT).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...   # the reference arithmetic on the host cores

A step = one pass of ``Decoder.__call__`` over one synthetic batch (config "Kokoro-82M full decoder + iSTFT
bf16, batch 64 synthetic sentences (~8 s each), 1 B200": B=64, T=320 frames = 8 s each, random-init weights).
With N>1 ranks (torchrun) every rank decodes its own batch (utterance sharding, weak scaling) and the
waveforms are gathered to rank 0 with NCCL inside the timed region.  Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from mlx_audio_b200 import synth  # noqa: E402

METRIC = "audio_seconds_generated_per_second_24khz_kokoro_decoder"
UNIT = "audio-s/s"
TC_NAMES = ("kk_conv1d_bf16_tc", "kk_conv1d_adain_bf16_tc")   # the tcgen05 implicit-GEMM convolutions


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(bf16_burst=d["bf16_tflops"], bf16_sustained=d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                    hbm=d["hbm_gbs"], source="measured (MEASURED_PEAKS.json)")
    return dict(bf16_burst=1590.0, bf16_sustained=1400.0, hbm=6650.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """Samples SM clocks / throttle reasons with nvidia-smi while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._read, daemon=True)
            self.thr.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return dict(sm_mhz=float(np.median(sm)) if sm else None, sm_max_mhz=max(mx) if mx else None,
                    reasons=sorted(reasons), samples=len(sm))


def run_reference(args, rank: int):
    """Reference arm: the reference's own arithmetic for this path on the host cores.  MLX (the library that
    carries the reference's kernels) cannot be installed here, so this times the CPU restatement in oracle/
    (kind "port") with every host thread, one utterance of the same shape per step."""
    if rank != 0:
        return
    import torch
    from oracle.istftnet import DecoderOracle
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    T = args.frames
    w = synth.synth_decoder_weights(seed=2000)
    orc = DecoderOracle(w)
    inp = synth.synth_decoder_inputs(1, T, seed=1000)
    call = lambda: orc(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], inp["noise"])
    for _ in range(max(1, min(args.warmup, 2))):
        call()
    steps = max(1, args.steps)
    t0 = time.perf_counter()
    for _ in range(steps):
        call()
    dt = (time.perf_counter() - t0) / steps
    val = (T * synth.SAMPLES_PER_FRAME / synth.SAMPLE_RATE) / dt
    sample = f"B=1 x T={T} ({T / 40:.1f} s of audio) per step, float32 torch-CPU/NumPy restatement, {cores} threads"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"Kokoro-82M ISTFTNet decoder + iSTFT, random-init, {sample}"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "MLX is not installable in this environment; the oracle port stands in for the MLX CPU path",
    }))


def run_ours(args, rank: int, world: int, local_rank: int):
    import torch
    import torch.distributed as dist
    from mlx_audio_b200 import _lib
    from mlx_audio_b200.tts.models.kokoro import build_decoder

    _lib.require_cuda()
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B, T = args.batch, args.frames
    S = T * synth.SAMPLES_PER_FRAME
    audio_s_per_step = B * S / synth.SAMPLE_RATE
    weights = synth.synth_decoder_weights(seed=2000)
    dec = build_decoder(weights, dtype=args.dtype, device=str(dev))
    dec.use_cuda_graph = bool(args.graph)
    inp = synth.synth_decoder_inputs(B, T, seed=1000 + 17 * rank, with_noise=False)
    host = {k: torch.from_numpy(v).pin_memory() for k, v in inp.items()}
    devt = {k: v.to(dev) for k, v in host.items()}
    gather_list = [torch.empty(B, 1, S, device=dev) for _ in range(world)] if (world > 1 and rank == 0) else None

    def step(src):
        y = dec(src["asr"], src["F0_curve"], src["N"], src["s"])
        if world > 1:  # the path's only exchange: waveform gather to rank 0
            dist.gather(y, gather_list, dst=0)
        return y

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for _ in range(max(3, args.warmup)):
        step(devt)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms_total = timed(lambda: step(devt), args.steps)
    clocks = sampler.stop() if rank == 0 else None

    # end to end through the public streaming API: every step copies its inputs from pinned host memory and its
    # waveforms back to pinned host memory; the copies of neighbouring steps overlap the forward (3 streams)
    from mlx_audio_b200.streaming import StreamingDecoder
    sd = StreamingDecoder(dec, depth=2, post=(lambda y: dist.gather(y, gather_list, dst=0)) if world > 1 else None)

    def e2e_run(steps):
        for wav in sd.run(host for _ in range(steps)):
            out_host = wav                   # pinned host tensor holding this step's (B, 1, 600 T) result
        return out_host
    e2e_run(2)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out_host = e2e_run(args.steps)           # returns after the last device->host copy has completed
    e1.record()
    barrier()
    e2e_t = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_ms = float(e2e_t.item())
    h2d = sum(v.numel() * v.element_size() for v in host.values())
    d2h = out_host.numel() * out_host.element_size()

    # per-kernel accounting on rank 0: CUDA events around every launch of one forward
    result = None
    if rank == 0:
        plan = dec._plan(B, T)
        recs = plan.run_timed(_lib.stream())
        recs = plan.run_timed(_lib.stream())
        by = {}
        tc_ms = tc_flops = 0.0
        for name, idx, ms in recs:
            a = by.setdefault(name, [0, 0.0])
            a[0] += 1; a[1] += ms
            if name in TC_NAMES:
                tc_ms += ms; tc_flops += plan.conv_flops(idx)
        tot = sum(v[1] for v in by.values())
        if args.ops_json:
            ops = []
            for name, idx, ms in recs:
                rec = {"op": name, "ms": ms}
                if idx in plan.conv_ops:
                    d = plan.conv_ops[idx]
                    fl = plan.conv_flops(idx)
                    rec.update(C_in=d.C_in, N=d.N, taps=d.taps, dil=d.dil, M=d.M, B=d.B, gflop=fl / 1e9,
                               tflops=fl / (ms / 1e3) / 1e12)
                ops.append(rec)
            json.dump(ops, open(args.ops_json, "w"), indent=0)
        peaks = load_peaks()
        ms_step = ms_total / args.steps
        value = world * audio_s_per_step / (ms_step / 1e3)
        e2e_val = world * audio_s_per_step / (e2e_ms / args.steps / 1e3)
        dom = "kk_conv1d_adain_bf16_tc" if (args.dtype == "bf16" and "kk_conv1d_adain_bf16_tc" in by) else (
            "kk_conv1d_bf16_tc" if args.dtype == "bf16" else "kk_conv1d_f32")
        peak = peaks["bf16_sustained"]
        # dominant kernel: algorithmic FLOPs of its launches / their CUDA-event durations (same stream, per launch)
        d_fl = sum(plan.conv_flops(i) for n, i, _ in recs if n == dom)
        d_ms, d_n = by[dom][1], by[dom][0]
        ach = d_fl / (d_ms / 1e3) / 1e12
        traffic, traffic_src = None, None
        tp = os.path.join(ROOT, "profiles", "r01_fused_traffic.json")
        if dom == "kk_conv1d_adain_bf16_tc" and os.path.exists(tp):
            tj = json.load(open(tp))
            traffic, traffic_src = tj["traffic_bytes_per_launch"], "profiles/r01_fused_traffic.json (ncu dram__bytes_read+write, mean per launch)"
        roof = {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                "traffic": traffic, "traffic_source": traffic_src,
                "kernel": {"kk_conv1d_adain_bf16_tc": "conv1d_adain_tc_kernel (fused AdaIN+Snake+conv, tcgen05)",
                           "kk_conv1d_bf16_tc": "conv1d_bf16_tc_persist_kernel (tcgen05)"}.get(dom, dom),
                "launches": d_n, "flop_per_launch": d_fl / d_n, "ms_per_launch": d_ms / d_n,
                "kernel_ms_per_step": d_ms, "share_of_step": d_ms / tot,
                "peak_source": peaks["source"] + ", sustained bf16 (kernel timed inside a long step)"}
        if args.dtype == "bf16" and tc_ms > 0:
            roof["all_tcgen05_kernels"] = {"achieved": tc_flops / (tc_ms / 1e3) / 1e12,
                                           "frac": tc_flops / (tc_ms / 1e3) / 1e12 / peak,
                                           "launches": sum(by[n][0] for n in TC_NAMES if n in by),
                                           "kernel_ms_per_step": tc_ms, "share_of_step": tc_ms / tot}
        else:
            roof["note"] = "fp32 CUDA-core mode; fraction still quoted against the bf16 tensor peak"
        end_to_end_tflops = value * synth.FLOP_PER_AUDIO_SECOND / 1e12 / world
        result = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
            "config": {"workload": f"Kokoro-82M ISTFTNet decoder + iSTFT, random-init weights, batch {B} x T={T} frames "
                                   f"({T / 40:.1f} s each) per GPU, {args.dtype}; utterance-sharded over {world} GPU(s)"
                                   + (", NCCL waveform gather to rank 0" if world > 1 else ""),
                       "batch_per_gpu": B, "frames": T, "audio_seconds_per_step": world * audio_s_per_step,
                       "l2": "per-step activations (~10 GB) far exceed the 126 MB L2; no explicit flush needed",
                       "cuda_graph": bool(args.graph)},
            "rtf": 1.0 / value,
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_ms / args.steps},
            "gpu_launches": plan.n_launches * args.steps,
            "launches_per_step": plan.n_launches,
            "roofline": roof,
            "decoder_tflops_per_gpu": end_to_end_tflops,
            "decoder_frac_of_bf16_peak": end_to_end_tflops / peaks["bf16_sustained"],
            "kernel_ms": {k: {"launches": v[0], "ms": round(v[1], 3)} for k, v in sorted(by.items(), key=lambda kv: -kv[1][1])},
            "clocks": clocks,
            "peak_memory_gb": torch.cuda.max_memory_allocated() / 1e9,
        }
    if world > 1:
        dist.barrier()
    # CPU baseline: rank 0, N=1 only, bounded sample of the same workload
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        import torch as _t
        from oracle.istftnet import DecoderOracle
        cores = os.cpu_count() or 1
        _t.set_num_threads(cores)
        Tc = min(T, 200)
        orc = DecoderOracle(weights)
        ci = synth.synth_decoder_inputs(1, Tc, seed=1000)
        orc(ci["asr"], ci["F0_curve"], ci["N"], ci["s"], ci["noise"])
        n_rep, t0 = 3, time.perf_counter()
        for _ in range(n_rep):
            orc(ci["asr"], ci["F0_curve"], ci["N"], ci["s"], ci["noise"])
        dt = (time.perf_counter() - t0) / n_rep
        result["cpu_baseline"] = {"value": (Tc / 40.0) / dt, "unit": UNIT, "cores": cores, "kind": "port",
                                  "sample": f"{n_rep} decodes of B=1 x T={Tc} ({Tc / 40:.1f} s audio), float32 oracle "
                                            f"(torch-CPU/NumPy restatement; MLX not installable), {cores} threads"}
    if rank == 0:
        print(json.dumps(result))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--frames", type=int, default=320)
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--graph", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--ops-json", default="", help="dump the per-launch event timings of one forward")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world == 1 and args.gpus > 1:
        # convenience: re-launch under torchrun
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
