cd /root/repo
python -m pytest tests -q -m gpu -x 2>&1 | tail -3
python bench.py --no-cpu-baseline > gpurun_out/ab.log 2>&1; python experiments/kms.py gpurun_out/ab.log | tail -2
python bench.py --no-cpu-baseline > gpurun_out/ab.log 2>&1; python experiments/kms.py gpurun_out/ab.log | tail -1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 291 -c 97 --csv --log-file gpurun_out/step_launches.csv \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_ll.log 2>&1
echo ncu rc $?
