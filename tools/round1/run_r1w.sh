#!/bin/bash
# developer (GPU box): tests, steady-state bench at 8192 and 16384 worlds, other models, parity report
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r1w_tests.txt
for nw in 8192 16384; do for f in all core; do
python bench.py --no-cpu --fields $f --nworld $nw --steps 400 --warmup 100 > gpurun_out/r1w_bench_${f}_${nw}.json 2> gpurun_out/r1w_bench_${f}_${nw}.err
done; done
for mdl in legs hand_sar finger elbow; do
python bench.py --no-cpu --model $mdl --fields all --steps 300 --warmup 100 > gpurun_out/r1w_bench_${mdl}.json 2> gpurun_out/r1w_bench_${mdl}.err
done
python tools/dev_parity.py > gpurun_out/r1w_parity.txt 2>&1
cat gpurun_out/r1w_tests.txt
python - <<'PY'
import json, glob
for p in sorted(glob.glob('gpurun_out/r1w_bench_*.json')):
    try:
        d = json.loads(open(p).read().strip().split('\n')[-1])
        print(p, 'ms/step %.3f value %.3e e2e %.3e frac %.4f bytes/world %d T%d x%d flagged %d' % (d['ms_per_step'], d['value'], d['e2e']['value'], d['roofline']['frac'], d['config']['bytes_per_world_step'], d['config']['lanes_per_world'], d['config']['step_ctas_per_sm'], d['config']['worlds_with_status_flags']))
    except Exception as e:
        print(p, 'FAILED', e, open(p.replace('.json', '.err')).read()[-400:])
PY
