#!/bin/bash
# developer (GPU box): larger batches per GPU, default build and two co-resident CTAs per SM (<= 72 registers)
out=gpurun_out/exp_nw.txt; : > $out
run() {  # label, team, nworld
  for f in core all; do
  MYO_TEAM=$2 python bench.py --no-cpu --fields $f --nworld $3 --steps 300 --warmup 100 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('[$1] T$2 nworld $3 $f: ms/step %.3f env-steps/s %.3e flagged %d' % (d['ms_per_step'], d['value'], d['config']['worlds_with_status_flags']))
    elif 'rror' in l: print(l.strip()[:200])
" >> $out
  done
}
run default 8 8192
run default 8 16384
run default 4 16384
run default 4 32768
(cd myosuite_mjx_b200/csrc; rm -f team8.o team4.o; make EXTRA="-DCTA_WARPS=14 -DCTA_MINBLOCKS=2" > /dev/null 2>&1)
run 14x2 8 16384
run 14x2 4 32768
(cd myosuite_mjx_b200/csrc; rm -f team8.o team4.o; make > /dev/null 2>&1)
cat $out
