#!/bin/bash
# Re-entry verification: GPU test suite, default bench, reference arm (short), config-2 STFT bench.
cd /root/repo
python -m pytest tests -q -m gpu -x 2>&1 | tail -5 > gpurun_out/verify_tests.txt
cat gpurun_out/verify_tests.txt
python bench.py > gpurun_out/verify_bench.json 2> gpurun_out/verify_bench.err; echo bench rc $?
python experiments/kms.py gpurun_out/verify_bench.json | tail -3
python experiments/bench_stft.py > gpurun_out/verify_stft.jsonl 2>&1; echo stft rc $?
tail -5 gpurun_out/verify_stft.jsonl
