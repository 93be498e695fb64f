#!/bin/bash
# A/B of the general FFT kernels inside ONE gpurun call: experiments/fft_ab.sh <libA.so> [reps]  (B = the product library)
cd "$(dirname "$0")/.."
A=$1; N=${2:-2}
mkdir -p gpurun_out
for i in $(seq $N); do
  KK_LIB=$PWD/$A timeout 300 python experiments/fft_ab.py gpurun_out/fft_ab_A$i.json > /dev/null 2> gpurun_out/fft_ab_A.err || tail -5 gpurun_out/fft_ab_A.err
  timeout 300 python experiments/fft_ab.py gpurun_out/fft_ab_B$i.json > /dev/null 2> gpurun_out/fft_ab_B.err || tail -5 gpurun_out/fft_ab_B.err
done
python - "$N" <<'PY'
import json, sys
n = int(sys.argv[1])
for i in range(1, n + 1):
    a = json.load(open(f"gpurun_out/fft_ab_A{i}.json")); b = json.load(open(f"gpurun_out/fft_ab_B{i}.json"))
    for k in a:
        if k == "lib": continue
        for kk in a[k]:
            if kk.endswith("_sha"):
                print(k, kk, "SAME" if a[k][kk] == b[k][kk] else f"DIFF {a[k][kk]} {b[k][kk]}")
            else:
                print(k, kk, f"A {a[k][kk]:.4f}  B {b[k][kk]:.4f}")
PY
