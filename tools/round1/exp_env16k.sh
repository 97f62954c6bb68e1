#!/bin/bash
# developer (GPU box): fused env steps (10 substeps per launch) at 16384 worlds under the kernel variants
out=gpurun_out/exp_env16k.txt; : > $out
for cfg in "default" "MYO_TEAM=8" "MYO_TEAM=4"; do
  echo "== $cfg" >> $out
  if [ "$cfg" = "default" ]; then python tools/env_bench.py 16384 60 2>&1 | grep "HandPose\|LegWalk" | cut -c1-150 >> $out
  else env $cfg python tools/env_bench.py 16384 60 2>&1 | grep "HandPose\|LegWalk" | cut -c1-150 >> $out; fi
done
for ns in 1 2 5 10; do
  for cfg in "MYO_CTAS=2 MYO_TEAM=8" "MYO_TEAM=8"; do
    env $cfg python bench.py --no-cpu --fields core --nworld 16384 --nsub $ns --steps 100 --warmup 30 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('[$cfg] nsub $ns: ms/launch %.3f env-steps/s %.3e' % (d['ms_per_step'], d['value']))
" >> $out
  done
done
cat $out
