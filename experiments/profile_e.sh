#!/bin/bash
cd /root/repo
WARM=0 REPS=1 ncu --set full --clock-control none --import-source on -k regex:"stft_fast_kernel|istft_stream_kernel" -c 3 -f -o /tmp/st \
  python experiments/bench_stft.py > gpurun_out/ncu_st.log 2>&1
echo ncu rc $?
ncu -i /tmp/st.ncu-rep --page raw --csv > gpurun_out/r01_stft_raw.csv
for i in 1 2 3; do
  ncu -i /tmp/st.ncu-rep --page details --launch-skip $((i-1)) --launch-count 1 > gpurun_out/r01_stft_details_$i.txt
  ncu -i /tmp/st.ncu-rep --page source --csv --launch-skip $((i-1)) --launch-count 1 > gpurun_out/r01_stft_src_$i.csv
done
ls -la gpurun_out | tail -8
