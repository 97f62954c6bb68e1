#!/bin/bash
# developer (GPU box): warps-per-tile sweep of one kernel of the pipeline; usage: tools/r2_sweep.sh VAR v1 v2 ... (200 steps after 100 warm-up)
var=$1; shift
for v in "$@"; do
env $var=$v python bench.py --no-cpu --steps 200 --warmup 100 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$var=$v: ms/step %.3f' % d['ms_per_step'])
    elif 'rror' in l: print(l.strip()[:200])
"
done
