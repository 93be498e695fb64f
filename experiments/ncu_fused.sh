#!/bin/bash
# ncu --set full (source-level) of fused launches 30..37 of a forward: C=128 k=3 (conv1, conv2) x3 iterations, then k=7
# conv1 / conv2.  usage: experiments/ncu_fused.sh [skip] [count] [tag]
cd "$(dirname "$0")/.."
SKIP=${1:-177}; COUNT=${2:-8}; TAG=${3:-r02_fused_k3k7}
CMD="python bench.py --steps 1 --warmup 3 --graph 0 --no-cpu-baseline --skip dsp,c4,c5,latency,fp32"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:conv1d_adain_tc_kernel -s $SKIP -c $COUNT -f \
    -o gpurun_out/$TAG $CMD > gpurun_out/ncu_run.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_run.log; ls -la gpurun_out/$TAG.ncu-rep
