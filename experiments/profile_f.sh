#!/bin/bash
cd /root/repo
ncu --set full --clock-control none --import-source on -k regex:conv1d_adain_tc_kernel -s 178 -c 1 -f -o /tmp/k3b \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_k3b.log 2>&1
echo ncu rc $?
ncu -i /tmp/k3b.ncu-rep --page details > gpurun_out/r01_k3b_details.txt
ncu -i /tmp/k3b.ncu-rep --page source --csv > gpurun_out/r01_k3b_src.csv
grep -n "Duration\|conv1d_adain" gpurun_out/r01_k3b_details.txt | head -5
