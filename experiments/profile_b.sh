#!/bin/bash
# call B: one ncu --set full capture of the fused conv kernel: launches of one forward (48) after 3 warm-up forwards
cd /root/repo
python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/b3.log 2>&1 || exit 4
ncu --set full --clock-control none --import-source on -k regex:conv1d_adain_tc_kernel -s 144 -c 48 -f -o /tmp/fu \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
echo ncu rc $?
ncu -i /tmp/fu.ncu-rep --page raw --csv > gpurun_out/r01_fused_final_raw.csv
for i in 47 48; do   # 1-based ids in the report: k=11 C=128 conv1-type / conv2-type (last resblock)
  ncu -i /tmp/fu.ncu-rep --page details --launch-skip $((i-1)) --launch-count 1 > gpurun_out/r01_fused_final_details_$i.txt
  ncu -i /tmp/fu.ncu-rep --page source --csv --launch-skip $((i-1)) --launch-count 1 > gpurun_out/r01_fused_final_src_$i.csv
done
ls -la gpurun_out /tmp/fu.ncu-rep
