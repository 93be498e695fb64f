cd /root/repo
export KK_RES16=${KK_RES16:-1} KK_KEEP_TAPS=${KK_KEEP_TAPS:-0}
python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/b3.log 2>&1 || exit 5
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:conv1d_adain_tc_kernel -s 147 -c 49 --csv --log-file gpurun_out/fused_launches.csv \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_ll.log 2>&1
echo ncu rc $?
