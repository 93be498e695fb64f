#!/bin/bash
cd /root/repo
ncu --set full --clock-control none --import-source on -k regex:conv1d_bf16_tc_persist -s 60 -c 20 -f -o /tmp/tc \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_tc.log 2>&1
echo ncu rc $?
ncu -i /tmp/tc.ncu-rep --page raw --csv > gpurun_out/r01_tc_raw.csv
for i in 17 18; do
  ncu -i /tmp/tc.ncu-rep --page details --launch-skip $((i-1)) --launch-count 1 > gpurun_out/r01_tc_details_$i.txt
  ncu -i /tmp/tc.ncu-rep --page source --csv --launch-skip $((i-1)) --launch-count 1 > gpurun_out/r01_tc_src_$i.csv
done
ls -la gpurun_out | tail -8
