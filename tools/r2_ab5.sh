python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python tools/dev_newton_timing.py 5 25 2>&1 | tail -14
python tools/dev_newton_timing.py 2>&1 | tail -14
for a in "--steps 20 --warmup 5" "--steps 200 --warmup 100"; do python bench.py --no-cpu --no-steady $a | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$a ms/step %.4f e2e %.3e' % (d['ms_per_step'], d['e2e']['value']))"; done
