#!/bin/bash
# source-level capture of the C=128 k=3 launches (conv1 = fused launch 30, conv2 = 31 of a forward)
cd /root/repo
ncu --set full --clock-control none --import-source on -k regex:conv1d_adain_tc_kernel -s 177 -c 2 -f -o /tmp/k3 \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_k3.log 2>&1
echo rc $?
cp /tmp/k3.ncu-rep gpurun_out/k3.ncu-rep
ls -la gpurun_out/k3.ncu-rep
