#!/bin/bash
# ncu --set full of the config-2 STFT / iSTFT / log-mel kernels (one launch each)
cd /root/repo
WARM=1 REPS=1 python experiments/bench_stft.py > gpurun_out/stft_plain.log 2>&1 || exit 1
WARM=0 REPS=1 ncu --set full --clock-control none --import-source on -k regex:'stft_fast_kernel|istft_stream_kernel' -c 3 -f -o /tmp/stft \
  python experiments/bench_stft.py > gpurun_out/ncu_stft.log 2>&1
echo ncu rc $?
ncu -i /tmp/stft.ncu-rep --page details > gpurun_out/stft_details.txt
ncu -i /tmp/stft.ncu-rep --page raw --csv > gpurun_out/stft_raw.csv
cp /tmp/stft.ncu-rep gpurun_out/stft.ncu-rep
