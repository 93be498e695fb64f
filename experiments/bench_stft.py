#!/usr/bin/env python
"""general STFT / iSTFT / fused log-mel on 256 signals x 720 000 samples (30 s at 24 kHz),
n_fft 1024, hop 256, Hann - timed with CUDA events through the C ABI (buffers resident), against the HBM floor of the
algorithmic bytes (stft 0.737 GB in + 2.955 GB out; istft the reverse; log-mel 0.737 GB in + 0.230 GB out).
Prints one JSON line per kernel; `python experiments/bench_stft.py > profiles/r01_stft_config2.jsonl`."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mlx_audio_b200 import _lib as L  # noqa: E402
from mlx_audio_b200 import utils as U  # noqa: E402


def main():
    L.require_cuda()
    lib = L.load()
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(
        os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0}
    B, n, n_fft, hop = 256, 720000, 1024, 256
    frames = 1 + n // hop
    bins = n_fft // 2 + 1
    g = torch.Generator(device="cuda").manual_seed(3001)
    x = 0.1 * torch.randn(B, n, device="cuda", generator=g)
    w = U.hanning(n_fft).contiguous()
    wp = U._resolve_window("hann", n_fft, periodic=True).contiguous()
    spec = torch.empty(B, frames, bins, 2, device="cuda")
    sig = torch.empty(B, (frames - 1) * hop, device="cuda")
    fb = U.mel_filters(24000, n_fft, 80).contiguous()
    mel = torch.empty(B, frames, 80, device="cuda")
    st = L.stream()
    calls = {
        "kk_stft_f32": (lambda: lib.kk_stft_f32(L.ptr(x), L.ptr(w), L.ptr(spec), B, n, n_fft, hop, 1, frames, st),
                        x.numel() * 4 + spec.numel() * 4),
        "kk_istft_f32": (lambda: lib.kk_istft_f32(L.ptr(spec), L.ptr(wp), L.ptr(sig), B, frames, n_fft, hop, n_fft // 2,
                                                  sig.shape[1], st),
                         spec.numel() * 4 + sig.numel() * 4),
        "kk_logmel_f32": (lambda: lib.kk_logmel_f32(L.ptr(x), L.ptr(w), L.ptr(fb), L.ptr(mel), B, n, n_fft, hop, 1, frames,
                                                    80, 1, 1, 1e-5, st),
                          x.numel() * 4 + mel.numel() * 4),
    }
    for name, (fn, nbytes) in calls.items():
        for _ in range(int(os.environ.get("WARM", "3"))):
            L.check(fn(), name)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = int(os.environ.get("REPS", "10"))
        e0.record()
        for _ in range(reps):
            L.check(fn(), name)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        gbs = nbytes / ms / 1e6
        print(json.dumps({"kernel": name, "workload": f"{B} x {n} samples, n_fft {n_fft}, hop {hop}", "ms": ms,
                          "algorithmic_gb": nbytes / 1e9, "achieved_gbs": gbs, "peak_gbs": peaks["hbm_gbs"],
                          "frac": gbs / peaks["hbm_gbs"], "audio_seconds_per_second": B * n / 24000 / (ms / 1e3),
                          "note": "working set 3.7 GB >> 126 MB L2"}))
    # round trip sanity: the reference normalises by sum(w), not sum(w^2) (utils.py:142-150), so
    # istft(stft(x)) = 0.75 x away from the edges for Hann with hop = n/4
    err = (sig[:, 2048:-2048] - 0.75 * x[:, 2048:sig.shape[1] - 2048]).abs().max().item()
    print(json.dumps({"check": "istft(stft(x)) - 0.75 x, max abs", "value": err}))


if __name__ == "__main__":
    main()
