#!/bin/bash
cd /root/repo
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "adain or fused" 2>&1 | tail -3
timeout 200 python experiments/fused_dbg.py 2>&1 | tail -8
python bench.py --no-cpu-baseline --ops-json gpurun_out/ops_new.json > gpurun_out/bench_new.log 2>&1; head -c 250 gpurun_out/bench_new.log
