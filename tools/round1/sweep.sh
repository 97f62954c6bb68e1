#!/bin/bash
# developer: bench at several batch sizes (core fields), prints ms/step and env-steps/s
for nw in "$@"; do
  python bench.py --steps 20 --warmup 3 --no-cpu --fields core --nworld $nw 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('nworld', $nw, 'ms/step %.3f' % d['ms_per_step'], 'env-steps/s %.3e' % d['value'], 'e2e %.3e' % d['e2e']['value'])
    elif 'Error' in l or 'error' in l: print(l.strip())
"
done
