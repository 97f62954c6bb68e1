#!/bin/bash
# developer (GPU box): GPU parity tests + steady-state bench (both field sets) -> gpurun_out/quick_*.txt
tag=${1:-q}
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/${tag}_tests.txt
for f in all core; do
python bench.py --no-cpu --fields $f --steps 400 --warmup 100 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$f: ms/step %.3f env-steps/s %.3e e2e %.3e flagged %d frac %.4f' % (d['ms_per_step'], d['value'], d['e2e']['value'], d['config']['worlds_with_status_flags'], d['roofline']['frac']))
    elif 'rror' in l: print(l.strip()[:300])
" >> gpurun_out/${tag}_bench.txt
done
cat gpurun_out/${tag}_tests.txt gpurun_out/${tag}_bench.txt
