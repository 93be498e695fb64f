cd /root/repo
for cfg in "KK_RES16=1 KK_KEEP_TAPS=0 KK_L2PF=1" "KK_RES16=1 KK_KEEP_TAPS=0 KK_L2PF=5" "KK_RES16=2 KK_KEEP_TAPS=0 KK_L2PF=5" "KK_RES16=0 KK_L2PF=5" "KK_RES16=1 KK_KEEP_TAPS=0 KK_L2PF=1" "KK_RES16=1 KK_KEEP_TAPS=0 KK_L2PF=5"; do
  env $cfg python bench.py --no-cpu-baseline > gpurun_out/ab.log 2>&1; echo "$cfg"; python experiments/kms.py gpurun_out/ab.log | tail -2 | cut -c1-200
done
