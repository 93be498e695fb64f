#!/bin/bash
cd /root/repo
unset KK_TC_BN; timeout 300 python bench.py --no-cpu-baseline --steps 3 --ops-json gpurun_out/ops_bn128.json > /dev/null 2>&1
KK_TC_BN=256 timeout 300 python bench.py --no-cpu-baseline --steps 3 --ops-json gpurun_out/ops_bn256.json > /dev/null 2>&1
ls -la gpurun_out/ops_bn*.json
