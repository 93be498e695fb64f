#!/bin/bash
cd /root/repo
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for cl in 0 1 3; do
  KK_FUSED_CL=$cl timeout 300 python bench.py --no-cpu-baseline --ops-json gpurun_out/ops_cl$cl.json > gpurun_out/bench_cl$cl.log 2>&1; echo "CL=$cl"; python experiments/kms.py gpurun_out/bench_cl$cl.log
done
