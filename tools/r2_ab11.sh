run() { tag=$1; shift; env "$@" python bench.py --no-cpu --no-steady --steps 300 --warmup 100 | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$tag $* steady-ish ms/step %.4f e2e %.3e' % (d['ms_per_step'], d['e2e']['value']))"
env "$@" python bench.py --no-cpu --no-steady --steps 20 --warmup 5 | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$tag $* short ms/step %.4f e2e %.3e' % (d['ms_per_step'], d['e2e']['value']))"; }
true
true
run base
for v in "$@"; do run $v MYO_LIB=$PWD/gpurun_variants/lib_$v.so; done
run base
