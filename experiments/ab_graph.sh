cd /root/repo
for i in 1 2; do for g in 0 1; do python bench.py --no-cpu-baseline --graph $g > gpurun_out/ab.log 2>&1; echo "graph=$g"; python experiments/kms.py gpurun_out/ab.log | tail -1; done; done
