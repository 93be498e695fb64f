#!/bin/bash
# Round-1 evidence of the final build, one gpurun call: default bench (with CPU baseline), reference arm, smoke,
# config-2 STFT bench, then ncu passes (each after its plain command exited 0): launch list of one step, --set full of
# the fused kernel's 49 launches of one forward, --set full of the config-2 STFT / iSTFT / log-mel kernels.
cd /root/repo
O=gpurun_out
python bench.py > $O/r01_bench_final.json 2> $O/r01_bench_final.err || exit 1
python bench.py --impl reference --steps 2 --warmup 1 > $O/r01_bench_reference_arm.json 2>> $O/r01_bench_final.err || exit 2
python -c "import __graft_entry__ as g; g.smoke()" > $O/r01_smoke.txt 2>&1 || exit 3
python experiments/bench_stft.py > $O/r01_stft_config2.jsonl 2>&1 || exit 4
python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $O/b3.log 2>&1 || exit 5
# launches of the last (timed) step: 4 steps x 97 launches, skip the first 3 steps
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 291 -c 97 --csv \
  --log-file $O/r01_launches_final.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $O/ncu_ll.log 2>&1
echo launch-list rc $?
ncu --set full --clock-control none --import-source on -k regex:conv1d_adain_tc_kernel -s 147 -c 49 -f -o /tmp/fu \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $O/ncu_full.log 2>&1
echo fused-full rc $?
ncu -i /tmp/fu.ncu-rep --page raw --csv > $O/r01_fused_final_raw.csv
ncu -i /tmp/fu.ncu-rep --page details --launch-skip 46 --launch-count 1 > $O/r01_fused_final_details_k11_conv1.txt
ncu -i /tmp/fu.ncu-rep --page details --launch-skip 47 --launch-count 1 > $O/r01_fused_final_details_k11_conv2.txt
WARM=0 REPS=1 ncu --set full --clock-control none --import-source on -k regex:'stft_fast_kernel|istft_stream_kernel' -c 3 -f -o /tmp/stft \
  python experiments/bench_stft.py > $O/ncu_stft.log 2>&1
echo stft-full rc $?
ncu -i /tmp/stft.ncu-rep --page details > $O/r01_stft_config2_ncu_details.txt
ncu -i /tmp/stft.ncu-rep --page raw --csv > $O/r01_stft_config2_ncu_raw.csv
python experiments/kms.py $O/r01_bench_final.json
tail -2 $O/r01_smoke.txt
cut -c1-150 $O/r01_stft_config2.jsonl
