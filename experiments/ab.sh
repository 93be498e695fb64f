#!/bin/bash
# usage: ab.sh "ENV_A" "ENV_B" [reps]   - alternate two environments inside one GPU session
cd /root/repo
A="$1"; B="$2"; N=${3:-2}
for i in $(seq $N); do
  for v in "$A" "$B"; do
    env $v python bench.py --no-cpu-baseline --steps 10 > gpurun_out/ab.log 2>&1; echo "[$v]"; python experiments/kms.py gpurun_out/ab.log | tail -1
  done
done
