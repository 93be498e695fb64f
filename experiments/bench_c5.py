#!/usr/bin/env python
"""long-form synthesis - 10 h of audio (36 000 s = 4 500 chunks of T = 320 asr frames, 8 s each)
streamed through the decoder from HOST buffers: every batch's inputs (asr / F0 / N / style) start in pinned host memory,
every waveform ends in pinned host memory (``mlx_audio_b200/streaming.py``: H2D of batch i+1 and D2H of batch i-1 under
the forward of batch i).  Chunks are dealt to the ranks of one box in contiguous runs (an audiobook's order is kept per
rank); there is no collective on the data path - each rank's audio lands in its own host buffer.  Whole-job
audio-seconds per second, wall time of the slowest rank (perf_counter around the stream, synchronised both sides, MAX
over ranks) - host copies included, that is the point of this config.

    python experiments/bench_c5.py                       # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 \
        experiments/bench_c5.py

Synthetic inputs: `--distinct` different pinned host batches per rank, cycled (generating 3 GB of distinct host noise
would only time the host RNG); random-init weights (seed 2000), the same recipe as bench.py."""
import argparse
import json
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mlx_audio_b200 import _lib, synth  # noqa: E402
from mlx_audio_b200.streaming import StreamingDecoder  # noqa: E402
from mlx_audio_b200.tts.models.kokoro import build_decoder  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--hours", type=float, default=10.0)
    ap.add_argument("--frames", type=int, default=320)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--distinct", type=int, default=4)
    ap.add_argument("--dtype", default="bf16")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    _lib.require_cuda()
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    T, B = args.frames, args.batch
    chunk_s = 600 * T / 24000.0
    chunks = int(round(args.hours * 3600 / chunk_s))
    lo, hi = rank * chunks // world, (rank + 1) * chunks // world      # contiguous run of chunks per rank
    sizes = [min(B, hi - k) for k in range(lo, hi, B)]
    dec = build_decoder(synth.synth_decoder_weights(seed=2000), dtype=args.dtype, device=str(dev))
    dec.reserve(B, T)
    host = []
    for d in range(args.distinct):
        inp = synth.synth_decoder_inputs(B, T, seed=5000 + 17 * rank + d, with_noise=False)
        host.append({k: torch.from_numpy(inp[k]).pin_memory() for k in ("asr", "F0_curve", "N", "s")})

    def batches():
        for i, n in enumerate(sizes):
            h = host[i % len(host)]
            yield h if n == B else {k: v[:n] for k, v in h.items()}

    def run_once():
        sd = StreamingDecoder(dec, depth=2)
        samples, peak = 0, 0.0
        for wav in sd.run(batches()):
            samples += wav.shape[0] * wav.shape[2]
            peak = max(peak, float(wav[0, 0, :1024].abs().max()))     # touch the host result
        return samples, peak

    # untimed warm-up: plans for the full and the ragged last batch, pinned staging slots
    sd = StreamingDecoder(dec, depth=2)
    for _ in sd.run(({k: v[:n] for k, v in host[0].items()} for n in sorted(set(sizes)))):
        pass
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    samples, peak = run_once()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    tt = torch.tensor([dt], device=dev, dtype=torch.float64)
    ss = torch.tensor([samples], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dist.all_reduce(ss, op=dist.ReduceOp.SUM)
    if rank == 0:
        audio_s = float(ss.item()) / 24000.0
        print(json.dumps({
            "config": "C5 long-form streaming", "n_gpus": world, "hours_of_audio": audio_s / 3600.0, "chunks": chunks,
            "frames_per_chunk": T, "batch": B, "batches_per_rank": len(sizes), "dtype": args.dtype,
            "seconds": float(tt.item()), "audio_seconds_per_second": audio_s / float(tt.item()),
            "rtf": float(tt.item()) / audio_s,
            "h2d_bytes_rank0": int(sum(n for n in sizes) * (512 * T + 4 * T + 128) * 4),
            "d2h_bytes_rank0": int(sum(n for n in sizes) * 600 * T * 4),
            "timing": "host wall clock around the whole stream (copies included), max over ranks",
            "data": f"synthetic, {args.distinct} distinct pinned host batches per rank cycled", "finite": peak == peak}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
