#!/bin/bash
# developer (GPU box): two co-resident 14-warp CTAs per SM (<= 72 registers) against the default one CTA per SM, T = 8 and 16
out=gpurun_out/exp_occ2.txt; : > $out
run() {  # $1 = label, $2 = MYO_TEAM or empty
  for f in core all; do
  MYO_TEAM=$2 python bench.py --no-cpu --fields $f --steps 400 --warmup 100 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('[$1] $f: ms/step %.3f flagged %d' % (d['ms_per_step'], d['config']['worlds_with_status_flags']))
    elif 'rror' in l: print(l.strip()[:200])
" >> $out
  done
}
build() { (cd myosuite_mjx_b200/csrc; rm -f team8.o team16.o; make EXTRA="$1" 2>&1 | grep -i "error\|spill" | head -5 >> ../../$out); }
run "default T8" 8
run "default T16 (two waves)" 16
build "-DCTA_WARPS=14 -DCTA_MINBLOCKS=2"
run "14x2 T16 (28 warps/SM)" 16
run "14x2 T8 (72-reg cap, one CTA/SM)" 8
build "-DCTA_WARPS=7 -DCTA_MINBLOCKS=4"
run "7x4 T16 (28 warps/SM in 4 CTAs)" 16
build ""
cat $out
