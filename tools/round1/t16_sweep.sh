#!/bin/bash
# developer (GPU box): TEAM 16 at the headline batch with single-wave CTA shapes -> gpurun_out/t16_sweep.txt
out=gpurun_out/t16_sweep.txt; : > $out
for cfg in "14 2" "28 1" "8 4"; do
  set -- $cfg
  (cd myosuite_mjx_b200/csrc; rm -f team16.o; make -j16 TEAMS="1 2 4 8 16" EXTRA="-DCTA_WARPS=$1 -DCTA_MINBLOCKS=$2" > /dev/null 2>&1)
  echo "== TEAM 16 CTA_WARPS=$1 CTA_MINBLOCKS=$2" >> $out
  for f in core all; do
  MYO_TEAM=16 python bench.py --no-cpu --fields $f --steps 400 --warmup 100 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('   $f: ms/step %.3f env-steps/s %.3e e2e %.3e' % (d['ms_per_step'], d['value'], d['e2e']['value']))
    elif 'rror' in l: print('   ', l.strip()[:200])
" >> $out
  done
done
cat $out
