# developer (GPU box): [TESTS=1] [TIMING=1] bash tools/r2_ab11.sh variant...   -> default build, each gpurun_variants/lib_<variant>.so, default build again
run() { tag=$1; shift; env "$@" python bench.py --no-cpu --no-steady --steps 300 --warmup 100 | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$tag $* steady-ish ms/step %.4f e2e %.3e' % (d['ms_per_step'], d['e2e']['value']))"
env "$@" python bench.py --no-cpu --no-steady --steps 20 --warmup 5 | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$tag $* short ms/step %.4f e2e %.3e' % (d['ms_per_step'], d['e2e']['value']))"; }
[ -n "$TESTS" ] && python -m pytest tests -m gpu -x -q 2>&1 | tail -2
[ -n "$TIMING" ] && python tools/dev_newton_timing.py | tail -16
run base
for v in "$@"; do run $v MYO_LIB=$PWD/gpurun_variants/lib_$v.so; done
[ $# -gt 0 ] && run base
