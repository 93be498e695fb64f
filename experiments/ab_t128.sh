#!/bin/bash
# C=128 stage of the fused kernel: 128-row tiles <1,128,2,SWAP> vs 256-row tiles <2,128,2,SWAP>, per launch (lab build)
cd "$(dirname "$0")/.."
export KK_LIB=$PWD/mlx_audio_b200/libkokoro_b200_tune.so
KK_FUSED_T128=31 timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_decoder.py -q -m gpu -x -k "adain or decoder" 2>&1 | tail -4
for sw in 0 31 0 31; do
  KK_FUSED_T128=$sw timeout 600 python bench.py --no-cpu-baseline --skip dsp,c4,c5,latency,fp32 --steps 8 --ops-json gpurun_out/ops_t128_${sw}.json > gpurun_out/ab_tmp.json 2> gpurun_out/ab_tmp.err || tail -3 gpurun_out/ab_tmp.err
  echo "== KK_FUSED_T128=$sw"; python experiments/kms.py gpurun_out/ab_tmp.json | tail -1
done
