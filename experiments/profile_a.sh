#!/bin/bash
# call A: default bench, reference arm, smoke, then the ncu launch list of the same bench command
set -x
cd /root/repo
python bench.py > gpurun_out/r01_bench.json 2> gpurun_out/r01_bench.err || exit 1
tail -c 600 gpurun_out/r01_bench.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r01_bench_ref.json 2>> gpurun_out/r01_bench.err || exit 2
cat gpurun_out/r01_bench_ref.json
python -c "import __graft_entry__ as g; g.smoke()" || exit 3
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/b2.log 2>&1 || exit 4
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r01_launches_final.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
echo ncu rc $?
