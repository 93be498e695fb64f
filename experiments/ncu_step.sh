#!/bin/bash
# Round-2 ncu evidence; ONE ncu pass per gpurun call (usage: ncu_step.sh list | metrics), after the same command exited 0
# without ncu:
#   (1) launch list of one step: every launch with its device time (cold-cache, serialised: compare SHARES);
#   (2) per-launch tensor-pipe / DRAM metrics of the 49 fused launches of one forward.
cd "$(dirname "$0")/.."
CMD="python bench.py --steps 1 --warmup 3 --graph 0 --parallel 0 --no-cpu-baseline --skip dsp,c4,c5,latency,fp32"   # --parallel 0: launches in plan order (the per-group labels below go by launch index)
MODE=${1:-list}
$CMD > gpurun_out/ncu_plain.log 2>&1 || exit 1
if [ "$MODE" = "list" ]; then
# bench.py runs 3 warm-up + 1 timed + 3 + 5 (graph A/B arm, eager here) forwards before the e2e part; one forward = 97 launches
ncu --metrics gpu__time_duration.sum --clock-control none -s 291 -c 97 --csv --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/ncu_l.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/r02_launches.csv
else
ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,dram__bytes_read.sum,dram__bytes_write.sum,sm__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,sm__cycles_elapsed.avg \
    --clock-control none -k regex:conv1d_adain_tc_kernel -s 147 -c 49 --csv --log-file gpurun_out/r02_fused_metrics.csv $CMD > gpurun_out/ncu_m.log 2>&1
echo "fused metrics rc=$?"; wc -l gpurun_out/r02_fused_metrics.csv
fi
