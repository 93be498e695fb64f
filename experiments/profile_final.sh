#!/bin/bash
# Final round-1 evidence: default bench (with CPU baseline), reference arm, smoke, config-2 STFT bench, then ONE ncu
# pass (launch list + full set of the fused kernel's launches of one forward).
cd /root/repo
python bench.py > gpurun_out/r01_bench_final.json 2> gpurun_out/r01_bench_final.err || exit 1
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r01_bench_reference_arm.json 2>> gpurun_out/r01_bench_final.err || exit 2
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r01_smoke.txt 2>&1 || exit 3
python experiments/bench_stft.py > gpurun_out/r01_stft_config2.jsonl 2>&1 || exit 4
python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/b3.log 2>&1 || exit 5
ncu --set full --clock-control none -k regex:conv1d_adain_tc_kernel -s 147 -c 49 -f -o /tmp/fu \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
echo ncu rc $?
ncu -i /tmp/fu.ncu-rep --page raw --csv > gpurun_out/r01_fused_final_raw.csv
ncu -i /tmp/fu.ncu-rep --page details --launch-skip 46 --launch-count 1 > gpurun_out/r01_fused_final_details_k11_conv1.txt
ncu -i /tmp/fu.ncu-rep --page details --launch-skip 47 --launch-count 1 > gpurun_out/r01_fused_final_details_k11_conv2.txt
python experiments/kms.py gpurun_out/r01_bench_final.json
cat gpurun_out/r01_smoke.txt | tail -2
