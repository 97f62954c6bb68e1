# developer (GPU box): bash tools/r2_env_ab.sh "ENV=1 ..." "ENV=2" ...   A/B of environment knobs on the default build
run() { env $1 python bench.py --no-cpu --no-steady --steps 300 --warmup 100 | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1 steady-ish ms/step %.4f e2e %.3e' % (d['ms_per_step'], d['e2e']['value']))"
env $1 python bench.py --no-cpu --no-steady --steps 20 --warmup 5 | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1 short ms/step %.4f e2e %.3e' % (d['ms_per_step'], d['e2e']['value']))"; }
for v in "$@"; do run "$v"; done
