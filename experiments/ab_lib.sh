#!/bin/bash
# A/B of two builds of the library inside ONE gpurun call (the power cap of the box drawn moves a step by a few %):
#   experiments/ab_lib.sh <libA.so|-> <libB.so|-> [reps]      ("-" = the product library)
# alternates the quick bench between them and prints per-kernel ms of every run.
cd "$(dirname "$0")/.."
A=$1; B=$2; N=${3:-2}
run() {
  local lib=$1
  if [ "$lib" = "-" ]; then unset KK_LIB; else export KK_LIB=$PWD/$lib; fi
  timeout 600 python bench.py --no-cpu-baseline --skip dsp,c4,c5,latency,fp32 --steps 8 > gpurun_out/ab_tmp.json 2> gpurun_out/ab_tmp.err || tail -3 gpurun_out/ab_tmp.err
  echo "== $lib"; python experiments/kms.py gpurun_out/ab_tmp.json | tail -1
  python - <<'PY'
import json
d=json.loads(open('gpurun_out/ab_tmp.json').read().strip().splitlines()[-1])
print({k: v['ms'] for k, v in list(d['kernel_ms'].items())[:3]})
PY
}
for i in $(seq $N); do run "$A"; run "$B"; done
