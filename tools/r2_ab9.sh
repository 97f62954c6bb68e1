run() { env "$@" python bench.py --no-cpu --no-steady --steps 300 --warmup 100 | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$* steady-ish ms/step %.4f e2e %.3e' % (d['ms_per_step'], d['e2e']['value']))"
env "$@" python bench.py --no-cpu --no-steady --steps 20 --warmup 5 | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$* short ms/step %.4f e2e %.3e' % (d['ms_per_step'], d['e2e']['value']))"; }
python tools/dev_timeline.py 300 10 | tail -15
run A=1
run MYO_NW_POSOUT=4
run MYO_NW_POSOUT=4 MYO_NW_OUT=4
run MYO_NW_TEN=6 MYO_NW_INE=6
run MYO_NW_TEN=12 MYO_NW_INE=12
run MYO_NW_COL=8 MYO_NW_MAKECON=8
