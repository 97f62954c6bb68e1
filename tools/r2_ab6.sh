run() { env "$@" python bench.py --no-cpu --no-steady --steps 200 --warmup 100 | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$* steady-ish ms/step %.4f e2e %.3e' % (d['ms_per_step'], d['e2e']['value']))"
env "$@" python bench.py --no-cpu --no-steady --steps 20 --warmup 5 | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$* short ms/step %.4f e2e %.3e' % (d['ms_per_step'], d['e2e']['value']))"; }
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python tools/dev_timeline.py 5 20 | tail -13
python tools/dev_timeline.py 300 20 | tail -13
run A=1
run MYO_PRIORITY=0
run MYO_NW_COL=8
run MYO_NW_COL=8 MYO_NW_MAKECON=8
run MYO_NW_INE=12
