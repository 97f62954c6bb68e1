#!/bin/bash
# developer (GPU box): GPU suite, then stream-ordered against pipelined wide outputs -> gpurun_out/exp_pipe.txt
out=gpurun_out/exp_pipe.txt; : > $out
timeout 600 python -m pytest tests -m gpu -q 2>&1 | tail -5 >> $out
for p in "" "--pipelined"; do
  timeout 300 python bench.py --no-cpu $p --steps 400 --warmup 100 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('[$p] ms/step %.3f value %.3e e2e %.3e frac %.4f launches %d flagged %d' % (d['ms_per_step'], d['value'], d['e2e']['value'], d['roofline']['frac'], d['gpu_launches'], d['config']['worlds_with_status_flags']))
    elif 'rror' in l: print(l.strip()[:200])
" >> $out
done
cat $out
