"""Why is the ragged corpus (bench c4) ~8 % below the uniform headline per audio second?  Per-launch CUDA events of one
ragged batch (B = 64, lengths ~ U[240, 400] padded to 400) against the uniform batch (B = 64 x T = 320), by kernel,
normalised to milliseconds per 512 s of audio."""
import collections, sys
import numpy as np, torch
sys.path.insert(0, ".")
from mlx_audio_b200 import _lib as L, synth
from mlx_audio_b200.tts.models.kokoro.istftnet import build_decoder

dev = "cuda"
dec = build_decoder(synth.synth_decoder_weights(seed=2000), dtype="bf16")
B = 64
rng = np.random.default_rng(7)
lens = rng.integers(240, 401, size=B)
res = {}
for tag, T, ln in (("uniform", 320, None), ("ragged", 400, lens)):
    g = torch.Generator(device=dev).manual_seed(1)
    asr = torch.randn(B, 512, T, device=dev, generator=g)
    F0 = torch.from_numpy(synth.synth_f0(B, 2 * T, seed=3)).to(dev)
    N = torch.randn(B, 2 * T, device=dev, generator=g)
    s = 0.5 * torch.randn(B, 128, device=dev, generator=g)
    kw = {} if ln is None else dict(lengths=torch.tensor(ln, dtype=torch.int32, device=dev))
    for _ in range(3):
        dec(asr, F0, N, s, borrow=True, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        dec(asr, F0, N, s, borrow=True, **kw)
    e1.record(); e1.synchronize()
    audio = (B * 320 if ln is None else int(ln.sum())) * 600 / 24000.0
    step = e0.elapsed_time(e1) / 5
    plan = dec._plan(B, T, ln is not None)
    dec.engine.activate(plan)
    recs = plan.run_timed(L.stream()); recs = plan.run_timed(L.stream())
    by = collections.OrderedDict()
    for name, idx, ms in recs:
        by[name] = by.get(name, 0.0) + ms
    k = 512.0 / audio
    res[tag] = (step * k, {n: v * k for n, v in by.items()})
    print(f"{tag}: {audio:.0f} s of audio, step {step:.2f} ms = {step * k:.2f} ms per 512 s; summed launches {sum(by.values()) * k:.2f}")
u, r = res["uniform"][1], res["ragged"][1]
for n in sorted(u, key=lambda n: -u[n]):
    print(f"  {n:30s} uniform {u[n]:7.3f}  ragged {r.get(n, 0):7.3f}  ({(r.get(n, 0) / u[n] - 1) * 100:+.1f} %)")
