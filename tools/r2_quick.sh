#!/bin/bash
# developer (GPU box): parity suite + quick steady-state benches of the phase pipeline; usage: tools/r2_quick.sh <tag> [extra env assignments...]
tag=${1:-r2q}; shift
out=gpurun_out
env "$@" timeout 900 python -m pytest tests -m gpu -q --tb=line 2>&1 | tail -25 > $out/${tag}_tests.txt
for f in all core; do
env "$@" timeout 600 python bench.py --no-cpu --fields $f --steps 400 --warmup 100 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$* $f: ms/step %.3f env-steps/s %.3e e2e %.3e flagged %d frac %.4f launches %d' % (d['ms_per_step'], d['value'], d['e2e']['value'], d['config']['worlds_with_status_flags'], d['roofline']['frac'], d['gpu_launches']))
    elif 'rror' in l: print(l.strip()[:300])
" >> $out/${tag}_bench.txt
done
cat $out/${tag}_tests.txt $out/${tag}_bench.txt
