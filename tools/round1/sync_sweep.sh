#!/bin/bash
# developer (GPU box): which CTA-wide phase barriers pay for themselves -> gpurun_out/sync_sweep.txt
out=gpurun_out/sync_sweep.txt; : > $out
run() {
  (cd myosuite_mjx_b200/csrc; rm -f team8.o; make EXTRA="$1" > /dev/null 2>&1)
  echo "== $1" >> $out
  for f in core all; do
  python bench.py --no-cpu --fields $f --steps 400 --warmup 100 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('   $f: ms/step %.3f env-steps/s %.3e' % (d['ms_per_step'], d['value']))
    elif 'rror' in l: print('   ', l.strip()[:200])
" >> $out
  done
}
run ""
run "-DNO_SYNC_AFTER_NEWTON"
run "-DNO_SYNC_AFTER_NEWTON -DNO_SYNC_AFTER_EULER"
(cd myosuite_mjx_b200/csrc; rm -f team8.o; make > /dev/null 2>&1)
cat $out
