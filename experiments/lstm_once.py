"""The native bi-LSTM of F0Ntrain at B=64 x L=320 (timing against cuDNN; one launch for an ncu capture)."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
from mlx_audio_b200 import synth
from mlx_audio_b200.tts.models.kokoro.modules import build_prosody_predictor
sys.path.insert(0, "tests")
from test_gpu_prosody import _torch_lstm

B, Lq = 64, 320
w = synth.synth_prosody_weights(seed=2100)
pp = build_prosody_predictor(w, dtype="fp32")
x = torch.as_tensor(synth.synth_prosody_inputs(B, Lq, seed=1)["x"]).cuda().transpose(1, 2).contiguous()
ref = _torch_lstm(w)


def timeit(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) / n


print(f"native  {timeit(lambda: pp.shared(x)):.3f} ms")
with torch.no_grad(), torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
    print(f"cuDNN (TF32 off) {timeit(lambda: ref(x)):.3f} ms")
with torch.no_grad(), torch.backends.cudnn.flags(enabled=True, allow_tf32=True):
    print(f"cuDNN (TF32 on)  {timeit(lambda: ref(x)):.3f} ms")
