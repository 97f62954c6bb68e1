for t in 4 8 16; do echo "== team $t"; MYO_TEAM=$t timeout 300 python tools/dev_parity.py hand 2>&1 | grep -E "bad|match|drift|FAILED" | cut -c1-200; MYO_TEAM=$t tools/sweep.sh 8192; done
