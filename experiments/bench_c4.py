#!/usr/bin/env python
"""a ragged corpus of 4096 utterances, T ~ U[240, 400] asr frames (seed 4001; --equal 320 for the
equal-length variant), decoded in same-length batches, sharded over the ranks of one box (batches balanced by
``sharding.assign_buckets``), waveforms gathered on rank 0 over NCCL.  Whole-job audio-seconds per second, timed on the
device (CUDA events, max over ranks) around decode + gather, after one untimed pass that builds the launch plans.

    python experiments/bench_c4.py                       # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        experiments/bench_c4.py

Synthetic inputs: asr / N drawn on the device per batch (seeded per batch), F0 contours from synth.synth_f0, random-init
weights (seed 2000) - the same recipe as bench.py."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mlx_audio_b200 import _lib, sharding, synth  # noqa: E402
from mlx_audio_b200.tts.models.kokoro import build_decoder  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--utterances", type=int, default=4096)
    ap.add_argument("--equal", type=int, default=0, help="all utterances of this many frames instead of U[240,400]")
    ap.add_argument("--max-batch", type=int, default=64)
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--dtype", default="bf16")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    _lib.require_cuda()
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    U = args.utterances
    frames = [args.equal] * U if args.equal else np.random.default_rng(4001).integers(240, 401, size=U).tolist()
    plan = sharding.assign_buckets(frames, world, args.max_batch)
    mine = plan[rank]
    dec = build_decoder(synth.synth_decoder_weights(seed=2000), dtype=args.dtype, device=str(dev))
    # inputs of this rank's batches, resident on the device before the timed region
    inputs = {}
    for t, ids in mine:
        g = torch.Generator(device=dev).manual_seed(1000 + ids[0])
        n = len(ids)
        inputs[(t, ids[0])] = dict(
            asr=torch.randn(n, 512, t, device=dev, generator=g),
            F0_curve=torch.from_numpy(synth.synth_f0(n, 2 * t, seed=1004 + ids[0])).to(dev),
            N=torch.randn(n, 2 * t, device=dev, generator=g),
            s=0.5 * torch.randn(n, 128, device=dev, generator=g))
    if mine:   # size the shared workspace once: largest batch x longest utterance this rank owns
        dec.reserve(max(len(ids) for _, ids in mine), max(t for t, _ in mine))

    def decode_batch(t, ids):
        x = inputs[(t, ids[0])]
        return dec(x["asr"], x["F0_curve"], x["N"], x["s"])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    t0 = time.perf_counter()
    out = sharding.decode_sharded(decode_batch, frames, max_batch=args.max_batch)   # untimed: builds the plans
    barrier()
    first_pass_s = time.perf_counter() - t0
    if rank == 0:
        assert len(out) == U and all(o.numel() == 600 * f for o, f in zip(out, frames))
        assert all(bool(torch.isfinite(o).all()) for o in out[:: max(1, U // 64)])
    del out
    times = []
    for _ in range(args.reps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = sharding.decode_sharded(decode_batch, frames, max_batch=args.max_batch)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        times.append(float(ms.item()))
        del out
    if rank == 0:
        audio_s = sum(frames) * synth.SAMPLES_PER_FRAME / synth.SAMPLE_RATE
        best = min(times)
        print(json.dumps({
            "metric": "audio_seconds_generated_per_second_24khz_kokoro_decoder", "value": audio_s / (best / 1e3),
            "unit": "audio-s/s", "n_gpus": world, "ms_total": best, "ms_all_reps": times, "scaling": "strong",
            "dtype": args.dtype, "data": "synthetic",
            "config": {"workload": f"C4: {U} utterances, " + (f"T={args.equal}" if args.equal else "T ~ U[240,400] (seed 4001)")
                       + f", same-length batches <= {args.max_batch}, sharded over {world} GPU(s), NCCL gather to rank 0",
                       "audio_seconds": audio_s, "batches_per_rank": [len(r) for r in plan],
                       "mean_batch": U / sum(len(r) for r in plan), "plans_rank0": len(dec.engine.plans)},
            "first_pass_s": first_pass_s, "peak_memory_gb": torch.cuda.max_memory_allocated() / 1e9}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
