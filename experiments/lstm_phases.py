"""kk_lstm_bidir_f32 alone (B=64 x L=320, H=256): time per launch and, on a -DKK_LSTM_DBG lab build, the clocks one thread
spends per phase of a step.  KK_LIB selects the library (experiments/ab_lib.sh convention)."""
import ctypes
import sys
import torch
sys.path.insert(0, ".")
from mlx_audio_b200 import _lib as L

B, Lq, H = 64, 320, 256
lib = L.load()
g = torch.Generator(device="cuda").manual_seed(1)
xw = torch.randn(B, Lq, 2, 4 * H, device="cuda", generator=g) * 0.5
whT = torch.randn(2, H, 4 * H, device="cuda", generator=g) * (1.0 / 16)
out = torch.empty(B, Lq, 2 * H, device="cuda")


def run():
    L.check(lib.kk_lstm_bidir_f32(L.ptr(xw), L.ptr(whT), L.ptr(out), B, Lq, H, None, L.stream()), "lstm")


import os
for _ in range(int(os.environ.get("WARM", "3"))):      # WARM=300: ~0.6 s of back-to-back launches, clocks fully ramped
    run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    run()
e1.record()
e1.synchronize()
print(f"kernel {e0.elapsed_time(e1) / 10:.3f} ms  ({e0.elapsed_time(e1) / 10 / Lq * 1e3:.2f} us per step)")
raw = ctypes.CDLL(lib._name)
if hasattr(raw, "kk_lstm_dbg_read"):
    buf = (ctypes.c_ulonglong * 8)()
    raw.kk_lstm_dbg_read(buf, 1)
    run()
    torch.cuda.synchronize()
    raw.kk_lstm_dbg_read(buf, 1)
    names = ["xw load issue", "matvec", "syncthreads", "gate + h sends", "wait for h(t-1)"]
    tot = sum(buf[:5])
    for n, c in zip(names, buf[:5]):
        print(f"  {n:18s} {c / Lq:8.0f} clk / step  {100 * c / tot:5.1f} %")
