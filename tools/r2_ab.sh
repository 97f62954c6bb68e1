#!/bin/bash
# developer (GPU box): A/B of library variants built by tools/build_variants.sh; usage: tools/r2_ab.sh tag1 tag2 ...   (300 steps after 100 warm-up, full mjData)
for tag in "$@"; do
MYO_LIB=$PWD/gpurun_variants/lib_$tag.so python bench.py --no-cpu --steps 300 --warmup 100 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$tag: ms/step %.3f flagged %d' % (d['ms_per_step'], d['config']['worlds_with_status_flags']))
    elif 'rror' in l: print(l.strip()[:200])
"
done
