#!/bin/bash
# developer (GPU box): one phase-locked 28-warp CTA per SM (<= 72 registers) against the default 14-warp CTA
out=gpurun_out/exp_cta28.txt; : > $out
run() {  # label, team, nworld, extra args
  for f in core all; do
  MYO_TEAM=$2 python bench.py --no-cpu --fields $f --nworld $3 $4 --steps 400 --warmup 100 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('[$1] T$2 nworld $3 $4 $f: ms/step %.3f env-steps/s %.3e flagged %d' % (d['ms_per_step'], d['value'], d['config']['worlds_with_status_flags']))
    elif 'rror' in l: print(l.strip()[:200])
" >> $out
  done
}
run "default 14x1" 8 8192
(cd myosuite_mjx_b200/csrc; rm -f team8.o team16.o; make EXTRA="-DCTA_WARPS=28 -DCTA_MINBLOCKS=1" 2>&1 | grep -i "error" >> ../../$out)
run "28x1" 16 8192
run "28x1" 8 16384
run "28x1" 8 16384 "--nsub 10"
(cd myosuite_mjx_b200/csrc; rm -f team8.o team16.o; make > /dev/null 2>&1)
cat $out
