run() { tag=$1; shift; env "$@" python bench.py --no-cpu --steps 200 --warmup 100 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$tag $*: ms/step %.3f' % d['ms_per_step'])
    elif 'rror' in l: print(l.strip()[:200])
"; }
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python tools/dev_newton_timing.py 2>&1 | tail -16
run cur
run nostream MYO_LIB=$PWD/gpurun_variants/lib_nostream.so
run cur
