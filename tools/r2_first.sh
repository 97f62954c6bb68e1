#!/bin/bash
# developer (GPU box): round-2 first light of the phase pipeline: parity suite on both engines, memcheck, quick benches
tag=${1:-r2a}
out=gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -60 > $out/${tag}_tests_phase.txt
MYO_ENGINE=mono timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -40 > $out/${tag}_tests_mono.txt
timeout 600 compute-sanitizer --tool memcheck --print-limit 20 python __graft_entry__.py smoke 2>&1 | tail -40 > $out/${tag}_memcheck.txt
for eng in phase mono; do
for f in all core; do
MYO_ENGINE=$eng timeout 600 python bench.py --no-cpu --fields $f --steps 400 --warmup 100 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$eng $f: ms/step %.3f env-steps/s %.3e e2e %.3e flagged %d frac %.4f launches %d' % (d['ms_per_step'], d['value'], d['e2e']['value'], d['config']['worlds_with_status_flags'], d['roofline']['frac'], d['gpu_launches']))
    elif 'rror' in l: print(l.strip()[:300])
" >> $out/${tag}_bench.txt
done
done
echo "== phase tests"; cat $out/${tag}_tests_phase.txt; echo "== mono tests"; tail -15 $out/${tag}_tests_mono.txt; echo "== memcheck"; tail -15 $out/${tag}_memcheck.txt; echo "== bench"; cat $out/${tag}_bench.txt
