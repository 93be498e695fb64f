#!/bin/bash
# Round-2 call 1: GPU test suite, smoke, the full default bench (all sub-objects), per-launch table, and the barrier-wait
# breakdown of the fused kernel on the lab (-DKK_FUSED_DBG) build.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/parity_measured.jsonl
timeout 900 python -m pytest tests -q -m gpu -x 2>&1 | tail -15 > gpurun_out/r02_tests.txt; cat gpurun_out/r02_tests.txt
timeout 300 python __graft_entry__.py --smoke 2>&1 | tail -4 > gpurun_out/r02_smoke.txt; cat gpurun_out/r02_smoke.txt
timeout 900 python bench.py --ops-json gpurun_out/r02_ops_a.json > gpurun_out/r02_bench_a.json 2> gpurun_out/r02_bench_a.err
echo "bench rc=$?"; head -c 600 gpurun_out/r02_bench_a.json; echo; tail -5 gpurun_out/r02_bench_a.err
DBG_SET=s16 timeout 300 python experiments/fused_dbg.py > gpurun_out/r02_fused_dbg.txt 2>&1; cat gpurun_out/r02_fused_dbg.txt
