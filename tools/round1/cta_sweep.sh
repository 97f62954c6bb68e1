#!/bin/bash
# developer (GPU box): CTA shape sweep of the step kernel at the headline batch (8192 worlds, TEAM 8) -> gpurun_out/cta_sweep.txt
out=gpurun_out/cta_sweep.txt; : > $out
for cfg in "8 2" "7 2" "14 1" "4 4" "5 3" "10 1"; do
  set -- $cfg
  (cd myosuite_mjx_b200/csrc; rm -f team8.o; make -j16 TEAMS="1 2 4 8 16" EXTRA="-DCTA_WARPS=$1 -DCTA_MINBLOCKS=$2" > /dev/null 2>&1)
  echo "== CTA_WARPS=$1 CTA_MINBLOCKS=$2" >> $out
  for f in core all; do
  python bench.py --no-cpu --fields $f --steps 400 --warmup 100 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('   $f: ms/step %.3f env-steps/s %.3e e2e %.3e' % (d['ms_per_step'], d['value'], d['e2e']['value']))
    elif 'rror' in l: print('   ', l.strip()[:200])
" >> $out
  done
done
cat $out
