#!/bin/bash
cd /root/repo
run() { echo "== $*"; env "$@" timeout 300 python bench.py --no-cpu-baseline > gpurun_out/ab_k.log 2>&1; python experiments/kms.py gpurun_out/ab_k.log | tail -1 | cut -c1-120; }
run KK_RES16=1
run KK_RES16=2
run KK_L2PF=4
run KK_L2PF=1
run KK_RES16=1
run KK_RES16=2
