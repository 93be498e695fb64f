"""Fuzz of the general STFT / iSTFT / log-mel kernels (utils.stft / istft / log_mel_spectrogram, batched and ragged) against
the CPU restatement of the reference (oracle/dsp.py): random n_fft (fast 1024 path and the generic mixed-radix path), hops,
padding modes, windows, batch sizes and per-signal lengths.  usage: python experiments/fuzz_dsp.py [cases] [seed]"""
import os
import sys
import traceback

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mlx_audio_b200 import utils as U      # noqa: E402
from oracle import dsp                     # noqa: E402

n_cases = int(sys.argv[1]) if len(sys.argv) > 1 else 60
r = np.random.Generator(np.random.PCG64(int(sys.argv[2]) if len(sys.argv) > 2 else 31))
fails = 0
n_fwd = n_inv = n_zero = 0
for i in range(n_cases):
    n_fft = int(r.choice([16, 20, 60, 64, 256, 400, 512, 1024, 1024, 1024, 2048]))
    hop = int(r.choice([n_fft // 4, n_fft // 2, n_fft, max(1, n_fft // 5), max(1, n_fft // 8)]))
    center = bool(r.integers(4))
    pad_mode = str(r.choice(["reflect", "constant"]))
    window = str(r.choice(["hann", "hamming", "blackman"]))
    B = int(r.integers(1, 6))
    n = int(r.integers(n_fft + 1, 6 * n_fft + 50))
    lens = [int(v) for v in r.integers(1, n + 1, size=B)]
    lens[int(r.integers(B))] = n
    x = r.standard_normal((B, n)).astype(np.float32)
    kw = dict(n_fft=n_fft, hop_length=hop, window=window, center=center, pad_mode=pad_mode)
    try:
        xs = torch.from_numpy(x).cuda()
        for b, l in enumerate(lens):
            xs[b, l:] = float("nan")
        S = U.stft(xs, lengths=lens, **kw)
        torch.cuda.synchronize()
        frs = []
        for b, l in enumerate(lens):
            try:
                ref = dsp.stft(x[b, :l], **kw)
            except (ValueError, IndexError):
                ref = None
            if ref is not None and center and pad_mode == "reflect" and l <= n_fft // 2:
                ref = None
            f = 0 if ref is None else ref.shape[0]
            frs.append(f)
            got = S[b].cpu().numpy()
            if f:
                err = np.abs(got[:f] - ref).max() / max(1.0, np.abs(ref).max())
                assert err < 2e-5, ("stft", b, l, err)
                n_fwd += 1
            else:
                n_zero += 1
            assert not np.abs(got[f:]).any(), ("stft tail", b, l)
        # inverse of the oracle's spectra (hop must divide into a valid synthesis: compare where the window sum is not tiny)
        if center and hop <= n_fft // 2 and window == "hann":
            Sx = torch.zeros(B, S.shape[1], n_fft // 2 + 1, dtype=torch.complex64)
            for b, l in enumerate(lens):
                if frs[b]:
                    Sx[b, :frs[b]] = torch.from_numpy(dsp.stft(x[b, :l], **kw))
            y = U.istft(Sx.transpose(1, 2).cuda(), hop_length=hop, win_length=n_fft, window=window, lengths=frs)
            for b in range(B):
                if frs[b] > 1:
                    ref = dsp.istft(Sx[b, :frs[b]].numpy().T, hop_length=hop, win_length=n_fft, window=window)
                    got = y[b].cpu().numpy()
                    err = np.abs(got[:ref.shape[0]] - ref).max() / max(1.0, np.abs(ref).max())
                    assert err < 1e-4, ("istft", b, frs[b], err)
                    n_inv += 1
                    assert not np.abs(got[ref.shape[0]:]).any(), ("istft tail", b)
    except Exception:
        fails += 1
        print(f"case {i} n_fft={n_fft} hop={hop} center={center} pad={pad_mode} win={window} B={B} n={n} lens={lens}: FAILED")
        print("   " + traceback.format_exc().strip().splitlines()[-1][:200])
print(f"fuzz_dsp: {n_cases} cases ({n_fwd} forward and {n_inv} inverse signals compared, {n_zero} too short to transform), {fails} failures")
sys.exit(1 if fails else 0)
