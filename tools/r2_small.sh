for cfg in "finger 4096" "elbow 65536" "finger 32768"; do set -- $cfg
for v in "A=1" "MYO_CONCURRENT=0" "MYO_OUT_LATE=0"; do
env $v python bench.py --no-cpu --no-steady --model $1 --nworld $2 --steps 300 --warmup 100 | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1 $2 $v ms/step %.4f value %.3e e2e %.3e' % (d['ms_per_step'], d['value'], d['e2e']['value']))"
done; done
