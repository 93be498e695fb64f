#!/bin/bash
# ncu --set full of one config-2 launch each of the STFT / iSTFT / log-mel kernels.  usage: experiments/ncu_fft.sh [tag]
cd "$(dirname "$0")/.."
TAG=${1:-r02_fft}
CMD="python experiments/fft_once.py"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'stft_fast_kernel|istft_stream_kernel|stft_r32_kernel|istft_r32_kernel' -c 3 -f \
    -o gpurun_out/$TAG $CMD > gpurun_out/ncu_run.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_run.log; ls -la gpurun_out/$TAG.ncu-rep
