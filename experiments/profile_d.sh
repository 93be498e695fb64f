#!/bin/bash
cd /root/repo
ncu --set full --clock-control none --import-source on -k regex:conv1d_adain_tc_kernel -s 174 -c 2 -f -o /tmp/k3 \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_k3.log 2>&1
echo ncu rc $?
for i in 1 2; do
  ncu -i /tmp/k3.ncu-rep --page details --launch-skip $((i-1)) --launch-count 1 > gpurun_out/r01_k3_details_$i.txt
  ncu -i /tmp/k3.ncu-rep --page source --csv --launch-skip $((i-1)) --launch-count 1 > gpurun_out/r01_k3_src_$i.csv
done
ls -la gpurun_out | tail -5
