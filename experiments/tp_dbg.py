"""Where does conv1d_tc_persist_kernel wait?  Walks one decoder forward (B=64 x T=320) op by op on a -DKK_TP_DBG lab
build (KK_DEFS=-DKK_TP_DBG KK_LAB_NAME=tpdbg) and prints, for the larger kk_conv1d_bf16_tc launches, the clocks one
thread per role of CTA 0 spent waiting, as a share of the kernel's clocks."""
import ctypes, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("KK_LIB", os.path.join(ROOT, "mlx_audio_b200", "libkokoro_b200_tpdbg.so"))
from mlx_audio_b200 import _lib as L, synth
from mlx_audio_b200.tts.models.kokoro.istftnet import build_decoder

B, T = 64, 320
dec = build_decoder(synth.synth_decoder_weights(seed=2000), dtype="bf16")
plan = dec._plan(B, T)
dec.engine.activate(plan)
raw = ctypes.CDLL(L.load()._name)
buf = (ctypes.c_ulonglong * 16)()
NAMES = ["TMA:empty", "MMA:tmem_empty", "MMA:full", "EPI:tmem_full", "EPI:res", "EPI:act_read", "EPI:res_tile_read"]
st = L.stream()
for rep in range(2):
    for a in plan.arenas:
        a.zero_()
    for i, (fn, args) in enumerate(plan.ops):
        if fn is None:
            continue
        is_tp = fn.__name__ == "kk_conv1d_bf16_tc"
        if is_tp:
            torch.cuda.synchronize(); raw.kk_tp_dbg_read(buf, 1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        L.check(fn(*args, st), fn.__name__)
        if is_tp:
            e1.record(); torch.cuda.synchronize(); raw.kk_tp_dbg_read(buf, 1)
            ms = e0.elapsed_time(e1)
            d = plan.conv_ops[i]
            if rep == 1 and ms > 0.2:
                tot = max(buf[7], 1)
                print(f"op {i}: C_in={d.C_in} taps={d.taps} N={d.N} up={d.up} M={d.M} res={bool(d.res)} out={bool(d.out)} act={bool(d.out_act)} "
                      f"stats={bool(d.stats)}: {ms:.3f} ms, {tot} clk; epilogue chunks of warp 2: {buf[8] / tot * 100:.0f} %;  "
                      + "  ".join(f"{n}={buf[j] / tot * 100:.0f}%" for j, n in enumerate(NAMES)))
                print("      inside a chunk (share of kernel): " + "  ".join(f"{n}={buf[j] / tot * 100:.0f}%" for j, n in
                      zip(range(10, 16), ["row+tmem_ld+res", "residual/bias/mask/scale", "act cvt+tile", "fp32 tile+store", "act store", "stats"])))
