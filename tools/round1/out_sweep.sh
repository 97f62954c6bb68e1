#!/bin/bash
# developer (GPU box): grid size sweep of the wide-output kernel -> gpurun_out/out_sweep.txt
out=gpurun_out/out_sweep.txt; : > $out
for nb in 592 1152 1184 2368 4736 9472; do
  MYO_OUT_BLOCKS=$nb python bench.py --no-cpu --steps 400 --warmup 100 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('blocks $nb: ms/step %.3f env-steps/s %.3e' % (d['ms_per_step'], d['value']))
    elif 'rror' in l: print('   ', l.strip()[:200])
" >> $out
done
cat $out
