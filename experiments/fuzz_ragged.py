#!/usr/bin/env python
"""Fuzz of Decoder.__call__(..., lengths=): random (B, T_pad, lengths) - including lengths whose row counts land exactly
on the 128 / 256-row tile boundaries of the tcgen05 kernels - every utterance of the ragged decode against decoding it
alone (<= 2e-5), padding silent, both modes, CUDA graph on for half of the cases.  python experiments/fuzz_ragged.py [n]"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mlx_audio_b200 import synth
from mlx_audio_b200.tts.models.kokoro import build_decoder

n_cases = int(sys.argv[1]) if len(sys.argv) > 1 else 24
w = synth.synth_decoder_weights(seed=2000)
decs = {"bf16": build_decoder(w, dtype="bf16"), "fp32": build_decoder(w, dtype="fp32")}
r = np.random.default_rng(20261105)
worst = 0.0
for case in range(n_cases):
    mode = "bf16" if case % 3 else "fp32"
    dec = decs[mode]
    B = int(r.integers(1, 12))
    T = int(r.choice([3, 16, 32, 47, 64, 96, 107, 128]))
    special = [t for t in (1, 2, 16, 32, 64, T) if t <= T]       # 120 t + 1 / 20 t rows around tile multiples
    lens = np.array([int(r.choice(special)) if r.random() < 0.4 else int(r.integers(1, T + 1)) for _ in range(B)], np.int32)
    inp = synth.synth_decoder_inputs(B, T, seed=9000 + case)
    for b, t in enumerate(lens):                                  # garbage in the padding
        inp["asr"][b, :, t:] = 1e3; inp["F0_curve"][b, 2 * t:] = 300.0; inp["N"][b, 2 * t:] = 50.0; inp["noise"][b, 600 * t:] = -7.0
    dec.use_cuda_graph = False
    y = dec(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], noise=inp["noise"], lengths=lens).cpu().numpy()
    assert np.isfinite(y).all(), (case, mode, B, T, lens)
    for b, t in enumerate(lens):
        one = dec(inp["asr"][b:b + 1, :, :t], inp["F0_curve"][b:b + 1, :2 * t], inp["N"][b:b + 1, :2 * t], inp["s"][b:b + 1],
                  noise=inp["noise"][b:b + 1, :600 * t]).cpu().numpy()
        d = float(np.abs(y[b, 0, :600 * t] - one[0, 0]).max())
        worst = max(worst, d)
        assert d <= 2e-5, (case, mode, B, T, lens.tolist(), b, d)
        assert not y[b, 0, 600 * t:].any(), (case, mode, b)
    if case % 2:                                                  # graph replay with device-resident lengths (Philox noise)
        dec.use_cuda_graph = True
        ld = torch.from_numpy(lens).cuda()
        g1 = dec(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], seed=case, lengths=ld).cpu().numpy()
        g2 = dec(inp["asr"], inp["F0_curve"], inp["N"], inp["s"], seed=case, lengths=ld).cpu().numpy()
        assert np.array_equal(g1, g2) and np.isfinite(g1).all()
        assert all(not g1[b, 0, 600 * t:].any() for b, t in enumerate(lens))
        dec.use_cuda_graph = False
    print(f"case {case:2d} {mode} B={B:2d} T={T:3d} lens={lens.tolist()} ok", flush=True)
print(f"fuzz_ragged: {n_cases} cases passed, worst |ragged - solo| = {worst:.2e}")
