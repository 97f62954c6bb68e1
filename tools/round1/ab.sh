#!/bin/bash
# developer (GPU box): same-box A/B of a team.cu build flag, alternating runs -> gpurun_out/ab.txt   usage: tools/ab.sh "-DFLAG"
out=gpurun_out/ab.txt; : > $out
for rep in 1 2; do
for fl in "" "$1"; do
  (cd myosuite_mjx_b200/csrc; rm -f team8.o; make EXTRA="$fl" > /dev/null 2>&1)
  for f in core all; do
  python bench.py --no-cpu --fields $f --steps 400 --warmup 100 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('[$fl] $f: ms/step %.3f' % d['ms_per_step'])
    elif 'rror' in l: print(l.strip()[:200])
" >> $out
  done
done
done
(cd myosuite_mjx_b200/csrc; rm -f team8.o; make > /dev/null 2>&1)
cat $out
