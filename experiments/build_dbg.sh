#!/bin/bash
# builds experiments/libkokoro_b200_dbg.so = the product library with the fused kernel's wait-time counters compiled in
set -e
cd "$(dirname "$0")/.."
objs=""
for f in common weights elementwise conv_f32 conv_tc conv_fused source istft_head fft; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -DKK_FUSED_DBG -c mlx_audio_b200/csrc/$f.cu -o /tmp/dbg_$f.o &
  objs="$objs /tmp/dbg_$f.o"
done
wait
nvcc -shared -o experiments/libkokoro_b200_dbg.so $objs -gencode arch=compute_100a,code=sm_100a -lcudart_static -Xcompiler -fPIC
