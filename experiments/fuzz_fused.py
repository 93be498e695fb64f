"""Fuzz of the fused AdaIN + Snake + conv kernel (kk_conv1d_adain_bf16_tc) on random shapes against the float64 reference
of tests/test_gpu_kernels.py::_fused_case: 128 / 256 channels, every kernel size and dilation of the generator, short and
long utterances, batches beyond 148 tiles, all input / residual variants.  usage: python experiments/fuzz_fused.py [cases] [seed]"""
import os
import sys
import traceback

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from mlx_audio_b200 import _lib as L      # noqa: E402
import test_gpu_kernels as K              # noqa: E402

n_cases = int(sys.argv[1]) if len(sys.argv) > 1 else 80
r = np.random.Generator(np.random.PCG64(int(sys.argv[2]) if len(sys.argv) > 2 else 21))
lib = L.load()
fails = 0
for i in range(n_cases):
    k = int(r.choice([1, 3, 7, 11]))
    dil = 1 if k == 1 else int(r.choice([1, 3, 5]))
    x_bf16 = bool(r.integers(2))
    case = dict(B=int(r.choice([1, 2, 3, 5, 40])), Lin=int(r.integers(max(8, (k - 1) * dil + 1), 900)), Cch=int(r.choice([128, 256])),
                k=k, dil=dil, x_bf16=x_bf16, act=str(r.choice(["snake", "snake", "leaky", "none"])),
                with_res=bool(r.integers(2)) and x_bf16, with_norm=bool(r.integers(4)), seed=i)
    if case["B"] == 40:
        case["Lin"] = min(case["Lin"], 300)
    junk = [torch.full((int(n),), float("nan"), device="cuda") for n in r.integers(1000, 2000000, size=4)]
    del junk
    try:
        K._fused_case(lib, **case)
    except L.KokoroB200Error as e:
        print(f"case {i} {case}: rejected: {str(e)[:90]}")
    except Exception:
        fails += 1
        print(f"case {i} {case}: FAILED")
        print("   " + traceback.format_exc().strip().splitlines()[-1][:200])
print(f"fuzz_fused: {n_cases} cases, {fails} failures")
sys.exit(1 if fails else 0)
