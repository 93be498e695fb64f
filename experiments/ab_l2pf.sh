#!/bin/bash
# A/B of the fused kernel's prefetch flags (KK_L2PF: bit 0 = L2 bulk prefetch of the next tile's input, bit 2 = residual
# prefetch off), alternating in one gpurun call because the box's power cap moves the step time by +-1 %.
cd /root/repo
for rep in 1 2 3; do for f in 5 4 0; do
  KK_L2PF=$f timeout 300 python bench.py --no-cpu-baseline --steps 20 > gpurun_out/ab_l.log 2>&1
  echo "flags=$f $(python experiments/kms.py gpurun_out/ab_l.log | tail -1 | cut -c1-75)"
done; done
