#!/bin/bash
# quick GPU check: test suite (optionally -k filter in $1) + short bench without the sub-benchmarks
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu -x ${1:+-k "$1"} 2>&1 | tail -25
timeout 600 python bench.py --no-cpu-baseline --skip dsp,c4,c5,latency,fp32 --ops-json gpurun_out/ops_quick.json > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err
echo "bench rc=$?"; python experiments/kms.py gpurun_out/bench_quick.json; tail -3 gpurun_out/bench_quick.err
