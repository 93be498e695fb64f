#!/bin/bash
# SM clock / power / throttle reasons sampled while kk_lstm_bidir_f32 runs back to back for ~4 s
cd "$(dirname "$0")/.."
nvidia-smi --query-gpu=clocks.sm,power.draw,clocks_throttle_reasons.active,temperature.gpu --format=csv,noheader -lms 250 > gpurun_out/lstm_clocks.csv &
SMI=$!
WARM=2000 timeout 100 python experiments/lstm_phases.py
kill $SMI
sort gpurun_out/lstm_clocks.csv | uniq -c | sort -rn | head -8
