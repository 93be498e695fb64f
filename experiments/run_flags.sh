for f in 1 3 5 9 15; do echo "flags=$f"; KK_L2PF=$f timeout 120 python experiments/fused_dbg.py 2>&1 | head -2; done
