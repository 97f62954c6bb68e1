run() { tag=$1; shift; env "$@" MYO_LIB=$PWD/gpurun_variants/lib_$tag.so python bench.py --no-cpu --steps 300 --warmup 100 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$tag $*: ms/step %.3f' % d['ms_per_step'])
    elif 'rror' in l: print(l.strip()[:200])
"; }
run base
run w16
run w16 MYO_NW_NEWTON=16
run w16 MYO_NW_NEWTON=16 MYO_NW_KIN=16 MYO_NW_TEN=16 MYO_NW_VEL=16 MYO_NW_COL=16 MYO_NW_INE=16 MYO_NW_FIN=16
run w16n1 MYO_NW_NEWTON=16
run w12 MYO_NW_NEWTON=12
run w12 MYO_NW_NEWTON=12 MYO_NW_KIN=12 MYO_NW_TEN=12 MYO_NW_VEL=12 MYO_NW_COL=12
