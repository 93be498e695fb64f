"""ProsodyPredictor.F0Ntrain at B=64 x L=320: end-to-end time (bf16 / fp32) and the per-launch breakdown of its plan."""
import sys, collections
import torch
sys.path.insert(0, ".")
from mlx_audio_b200 import _lib as L, synth
from mlx_audio_b200.tts.models.kokoro.modules import build_prosody_predictor

B, Lq = 64, 320
w = synth.synth_prosody_weights(seed=2100)
inp = synth.synth_prosody_inputs(B, Lq, seed=1)
x = torch.as_tensor(inp["x"]).cuda()
s = torch.as_tensor(inp["s"]).cuda()
for dt in ("bf16", "fp32"):
    pp = build_prosody_predictor(w, dtype=dt)
    for _ in range(3):
        pp.F0Ntrain(x, s, borrow=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        pp.F0Ntrain(x, s, borrow=True)
    e1.record(); e1.synchronize()
    print(f"{dt}: F0Ntrain {e0.elapsed_time(e1) / 10:.3f} ms per call (LSTM included, heads on two streams)")
    pp.parallel_heads = False
    for _ in range(3):
        pp.F0Ntrain(x, s, borrow=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(10):
        pp.F0Ntrain(x, s, borrow=True)
    e1.record(); e1.synchronize()
    print(f"{dt}: F0Ntrain {e0.elapsed_time(e1) / 10:.3f} ms per call (one stream)")
    plan = pp.engine.plans[(B, Lq)]
    recs = plan.run_timed(L.stream()); recs = plan.run_timed(L.stream())
    by = collections.OrderedDict()
    for name, idx, ms in recs:
        a = by.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += ms
    tot = sum(v[1] for v in by.values())
    print(f"   plan: {len(recs)} launches, {tot:.3f} ms summed")
    for n, (c, ms) in sorted(by.items(), key=lambda kv: -kv[1][1]):
        print(f"     {n:32s} {c:3d}  {ms:.3f} ms")
    for name, idx, ms in recs:
        if ms > 0.08:
            d = plan.conv_ops.get(idx)
            print(f"       op {idx} {name} {ms:.3f} ms" + (f"  C_in={d.C_in} N={d.N} taps={d.taps} M={d.M}" if d is not None else ""))
