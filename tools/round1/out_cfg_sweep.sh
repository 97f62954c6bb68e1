#!/bin/bash
# developer (GPU box): wide-output kernel build variants -> gpurun_out/out_cfg_sweep.txt
out=gpurun_out/out_cfg_sweep.txt; : > $out
run() {
  (cd myosuite_mjx_b200/csrc; rm -f out.o; make EXTRA="$1" > /dev/null 2>&1)
  echo "== $1 MYO_OUT_BLOCKS=$2" >> $out
  MYO_OUT_BLOCKS=$2 python bench.py --no-cpu --steps 400 --warmup 100 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('   all: ms/step %.3f env-steps/s %.3e' % (d['ms_per_step'], d['value']))
    elif 'rror' in l: print('   ', l.strip()[:200])
" >> $out
}
run "" 0
run "-DOUT_STCS" 0
run "-DOUT_THREADS=256 -DOUT_MINBLOCKS=4" 4736
run "-DOUT_THREADS=256 -DOUT_MINBLOCKS=4 -DOUT_STCS" 4736
run "-DOUT_THREADS=128 -DOUT_MINBLOCKS=12" 14208
run "-DOUT_THREADS=128 -DOUT_MINBLOCKS=12 -DOUT_STCS" 14208
run "-DOUT_THREADS=512 -DOUT_MINBLOCKS=2 -DOUT_STCS" 2368
run "-DOUT_THREADS=64 -DOUT_MINBLOCKS=16 -DOUT_STCS" 18944
(cd myosuite_mjx_b200/csrc; rm -f out.o; make > /dev/null 2>&1)
cat $out
