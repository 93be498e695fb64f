#!/bin/bash
cd /root/repo
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python bench.py --no-cpu-baseline --ops-json gpurun_out/ops_new.json > gpurun_out/bench_new.log 2>&1; head -c 250 gpurun_out/bench_new.log; echo
KK_TC_DIRECT=0 python bench.py --no-cpu-baseline --ops-json gpurun_out/ops_old.json > gpurun_out/bench_old.log 2>&1; head -c 250 gpurun_out/bench_old.log
