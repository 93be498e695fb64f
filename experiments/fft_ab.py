"""A/B of two builds of the general STFT / iSTFT / log-mel kernels (KK_LIB selects the library): SHA-256 of the outputs on
seeded inputs at several shapes (bit equality between builds = same hashes) and the config-2 timings (median of 20).
Usage: KK_LIB=... python experiments/fft_ab.py out.json"""
import hashlib
import json
import sys

import torch

sys.path.insert(0, ".")
from mlx_audio_b200 import _lib as L   # noqa: E402
from mlx_audio_b200 import utils as U  # noqa: E402


def run(lib, B, n, n_fft=1024, hop=256, time_it=False):
    dev = torch.device("cuda:0")
    frames, bins = 1 + n // hop, n_fft // 2 + 1
    g = torch.Generator(device=dev).manual_seed(3001 + B + n)
    x = 0.1 * torch.randn(B, n, device=dev, generator=g)
    w = U.hanning(n_fft).contiguous()
    wp = U._resolve_window("hann", n_fft, periodic=True).contiguous()
    spec = torch.empty(B, frames, bins, 2, device=dev)
    sig = torch.empty(B, (frames - 1) * hop, device=dev)
    fb = U.mel_filters(24000, n_fft, 80).contiguous()
    mel = torch.empty(B, frames, 80, device=dev)
    st = L.stream()
    calls = {
        "stft": lambda: lib.kk_stft_f32(L.ptr(x), L.ptr(w), L.ptr(spec), B, n, n_fft, hop, 1, frames, None, st),
        "istft": lambda: lib.kk_istft_f32(L.ptr(spec), L.ptr(wp), L.ptr(sig), B, frames, n_fft, hop, n_fft // 2, sig.shape[1], None, 0, st),
        "logmel": lambda: lib.kk_logmel_f32(L.ptr(x), L.ptr(w), L.ptr(fb), L.ptr(mel), B, n, n_fft, hop, 1, frames, 80, 1, 1, 1e-5, None, st),
    }
    res = {}
    for key, fn in calls.items():
        for _ in range(3 if time_it else 1):
            L.check(fn(), key)
        torch.cuda.synchronize()
        if time_it:
            ts = []
            for _ in range(20):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                L.check(fn(), key)
                e1.record()
                e1.synchronize()
                ts.append(e0.elapsed_time(e1))
            res[key + "_ms"] = sorted(ts)[len(ts) // 2]
    for key, t in (("stft", spec), ("istft", sig), ("logmel", mel)):
        res[key + "_sha"] = hashlib.sha256(t.cpu().numpy().tobytes()).hexdigest()[:16]
    return res


def main():
    lib = L.load()
    out = {"lib": L.LIB_PATH}
    for B, n in ((1, 24000), (3, 50000), (7, 131072 + 77), (256, 720000)):
        out[f"B{B}_n{n}"] = run(lib, B, n, time_it=(B == 256))
    json.dump(out, open(sys.argv[1], "w"), indent=1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
