#!/bin/bash
# per-launch comparison of the fused kernel's SWAP (weights = M operand) and non-SWAP forms on the 128-channel stage
cd "$(dirname "$0")/.."
export KK_LIB=$PWD/mlx_audio_b200/libkokoro_b200_tune.so
for sw in 1 0 1 0; do
  KK_FUSED_SWAP=$sw timeout 600 python bench.py --no-cpu-baseline --skip dsp,c4,c5,latency,fp32 --steps 8 --ops-json gpurun_out/ops_swap${sw}.json > gpurun_out/ab_tmp.json 2> gpurun_out/ab_tmp.err || tail -3 gpurun_out/ab_tmp.err
  echo "== KK_FUSED_SWAP=$sw"; python experiments/kms.py gpurun_out/ab_tmp.json | tail -1
done
