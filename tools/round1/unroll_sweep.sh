#!/bin/bash
# developer (GPU box): TFOR unroll factor sweep -> gpurun_out/unroll_sweep.txt
out=gpurun_out/unroll_sweep.txt; : > $out
for u in 1 2 4; do
  (cd myosuite_mjx_b200/csrc; rm -f team8.o; make -j16 EXTRA="-DTFOR_UNROLL=$u" > /dev/null 2>&1)
  echo "== TFOR_UNROLL=$u" >> $out
  for f in core all; do
  python bench.py --no-cpu --fields $f --steps 400 --warmup 100 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('   $f: ms/step %.3f env-steps/s %.3e' % (d['ms_per_step'], d['value']))
    elif 'rror' in l: print('   ', l.strip()[:200])
" >> $out
  done
done
cat $out
