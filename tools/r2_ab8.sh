run() { env "$@" python bench.py --no-cpu --no-steady --steps 300 --warmup 100 | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$* steady-ish ms/step %.4f e2e %.3e' % (d['ms_per_step'], d['e2e']['value']))"
env "$@" python bench.py --no-cpu --no-steady --steps 20 --warmup 5 | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$* short ms/step %.4f e2e %.3e' % (d['ms_per_step'], d['e2e']['value']))"; }
run MYO_DELAY=1
run MYO_DELAY=0
run MYO_DELAY=0 MYO_CONOUT_LATE=1
run MYO_DELAY=1 MYO_CONOUT_LATE=1
run MYO_DELAY=0 MYO_PRIORITY=0
run MYO_DELAY=0 MYO_PRIORITY=0 MYO_CONOUT_LATE=1
run MYO_CONCURRENT=0
run MYO_DELAY=0 MYO_NW_INE=12 MYO_NW_TEN=12
run MYO_DELAY=0 MYO_NW_MAKECON=8 MYO_NW_VEL=8
