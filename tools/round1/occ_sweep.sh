#!/bin/bash
# developer (GPU box): rebuild the step kernel at several (CTA_WARPS, CTA_MINBLOCKS) register budgets and time the bench at
# several team widths / batch sizes.  usage: tools/occ_sweep.sh  -> gpurun_out/occ_sweep.txt
out=gpurun_out/occ_sweep.txt; : > $out
run() {  # team nworld fields
  MYO_TEAM=$1 python bench.py --no-cpu --fields $3 --nworld $2 --steps 200 --warmup 20 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('   team $1 nworld $2 $3: ms/step %.3f env-steps/s %.3e e2e %.3e' % (d['ms_per_step'], d['value'], d['e2e']['value']))
    elif 'rror' in l: print('   ', l.strip()[:200])
" >> $out
}
for cfg in "8 2" "8 4" "4 8" "16 2" "8 3"; do
  set -- $cfg
  (cd myosuite_mjx_b200/csrc; rm -f team*.o; make -j16 TEAMS="1 2 4 8 16" EXTRA="-DCTA_WARPS=$1 -DCTA_MINBLOCKS=$2" > /dev/null 2>&1)
  echo "== CTA_WARPS=$1 CTA_MINBLOCKS=$2" >> $out
  for t in 8 16; do run $t 8192 core; run $t 8192 all; done
  run 8 16384 core; run 16 16384 core; run 4 16384 core; run 4 32768 core
done
cat $out
