"""Fuzz of the three convolution entry points (kk_conv1d_f32, kk_conv1d_bf16_tc, kk_conv1d_tf32_tc) on random shapes against
the float64 reference of tests/test_gpu_kernels.py::_conv_case: channel counts that are / are not multiples of the tile
widths, rows that are not multiples of 32, plain and polyphase (transposed) geometry, residuals, statistics, dirty
memory between calls.  usage: python experiments/fuzz_conv.py [cases] [seed]"""
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

n_cases = int(sys.argv[1]) if len(sys.argv) > 1 else 150
r = np.random.Generator(np.random.PCG64(int(sys.argv[2]) if len(sys.argv) > 2 else 11))
lib = L.load()
fails = 0
for i in range(n_cases):
    kind = ["f32", "bf16", "tf32x3"][int(r.integers(3))]
    up = int(r.choice([0, 0, 0, 2, 4, 6, 10]))
    cout_pool = [32, 64, 96, 128, 160, 192, 256, 320] if (up and kind != "f32") else [22, 32, 48, 64, 96, 128, 130, 160, 192, 256, 320]
    case = dict(B=int(r.integers(1, 4)), Lin=int(r.integers(5, 400)), Cin=int(r.choice([24, 64, 70, 96, 128, 144, 256])),
                Cout=int(r.choice(cout_pool)), k=0 if up else int(r.choice([1, 3, 7, 11])), dil=1 if up else int(r.choice([1, 3, 5])),
                transposed_up=up, with_res=bool(r.integers(2)), left_pad=int(r.integers(2)) if up else 0, seed=i)
    if not up and case["Lin"] <= (case["k"] - 1) * case["dil"]:
        case["Lin"] += (case["k"] - 1) * case["dil"]
    junk = [torch.full((int(n),), float("nan"), device="cuda") for n in r.integers(1000, 2000000, size=4)]
    del junk
    try:
        if kind == "bf16":
            K._conv_case(lib, True, **case)
        elif kind == "tf32x3":
            K._conv_case(lib, False, tf32_terms=3, **case)
        else:
            K._conv_case(lib, False, **case)
    except L.KokoroB200Error as e:          # a shape the entry point rejects: fine, as long as it says so
        print(f"case {i} {kind} {case}: rejected: {str(e)[:90]}")
    except Exception:
        fails += 1
        print(f"case {i} {kind} {case}: FAILED")
        print("   " + traceback.format_exc().strip().splitlines()[-1][:200])
print(f"fuzz_conv: {n_cases} cases, {fails} failures")
sys.exit(1 if fails else 0)
