"""Random shapes / ragged lengths for kk_lstm_bidir_f32 against torch.nn.LSTM (cuDNN, TF32 off) run per sequence.
Covers every hidden size, one and several waves of clusters, all three sequences-per-cluster variants, zero lengths."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
from mlx_audio_b200 import _lib as L

lib = L.load()
rng = np.random.default_rng(int(sys.argv[1]) if len(sys.argv) > 1 else 0)
n_cases = int(sys.argv[2]) if len(sys.argv) > 2 else 60
worst = 0.0
for case in range(n_cases):
    H = int(rng.choice([64, 128, 192, 256]))
    B = int(rng.choice([1, 2, 7, 8, 9, 15, 16, 17, 40, 56, 57, 64, 70, 71, 90, 91, 120, 130]))
    Lq = int(rng.integers(1, 41))
    D = int(rng.integers(1, 40))
    ragged = bool(rng.integers(0, 2))
    torch.manual_seed(case)
    m = torch.nn.LSTM(D, H, batch_first=True, bidirectional=True).cuda().eval()
    x = torch.randn(B, Lq, D, device="cuda")
    lens = rng.integers(0, Lq + 1, size=B) if ragged else np.full(B, Lq)
    if ragged and case % 3 == 0:
        lens[rng.integers(0, B)] = Lq
    xw = torch.cat([(x.double() @ getattr(m, f"weight_ih_l0{s}").double().t() + getattr(m, f"bias_ih_l0{s}").double()
                     + getattr(m, f"bias_hh_l0{s}").double()) for s in ("", "_reverse")], dim=2).float().contiguous()
    for b in range(B):
        xw[b, lens[b]:] = float("nan")                      # nothing past a sequence's end may be read into its result
    whT = torch.stack([m.weight_hh_l0.detach().t(), m.weight_hh_l0_reverse.detach().t()]).contiguous()
    out = torch.full((B, Lq, 2 * H), float("nan"), device="cuda")
    lt = torch.as_tensor(lens, dtype=torch.int32, device="cuda")
    L.check(lib.kk_lstm_bidir_f32(L.ptr(xw), L.ptr(whT), L.ptr(out), B, Lq, H, L.ptr(lt) if ragged else None, L.stream()), "lstm")
    torch.cuda.synchronize()
    assert torch.isfinite(out).all(), (case, "non-finite output")
    with torch.no_grad(), torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
        for b in range(B):
            n = int(lens[b])
            if n:
                ref = m(x[b:b + 1, :n])[0][0]
                err = float((out[b, :n] - ref).abs().max())
                worst = max(worst, err)
                assert err <= 2e-5, (case, H, B, Lq, b, n, err)
            assert float(out[b, n:].abs().max()) == 0.0 if n < Lq else True, (case, b, "rows past the end must be zero")
print(f"{n_cases} cases, worst |err| {worst:.2e}: ok")
