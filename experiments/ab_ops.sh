#!/bin/bash
# per-launch comparison of two builds inside one gpurun call: experiments/ab_ops.sh <libA.so> (B = the product library)
cd "$(dirname "$0")/.."
A=$1
for tag in A B A B; do
  if [ "$tag" = "A" ]; then export KK_LIB=$PWD/$A; else unset KK_LIB; fi
  timeout 600 python bench.py --no-cpu-baseline --skip dsp,c4,c5,latency,fp32 --steps 6 --graph 0 --ops-json gpurun_out/ops_$tag.json > /dev/null 2> gpurun_out/ab_ops.err || tail -3 gpurun_out/ab_ops.err
  [ -f gpurun_out/ops_${tag}1.json ] && mv gpurun_out/ops_$tag.json gpurun_out/ops_${tag}2.json || mv gpurun_out/ops_$tag.json gpurun_out/ops_${tag}1.json
done
python - <<'PY'
import json
def load(t): return [json.load(open(f"gpurun_out/ops_{t}{i}.json")) for i in (1, 2)]
A, B = load("A"), load("B")
n = len(A[0])
tot = [0, 0]
for i in range(n):
    a = min(x[i]["ms"] for x in A); b = min(x[i]["ms"] for x in B)
    tot[0] += a; tot[1] += b
    o = A[0][i]
    if abs(a - b) > 0.01:
        print(i, o["op"], f"C={o.get('C_in')} taps={o.get('taps')} res={o.get('res')} out={o.get('out')} add={o.get('out_add')}  A {a:.3f}  B {b:.3f}  {b / a:.2f}x")
print("total A %.2f  B %.2f" % tuple(tot))
PY
