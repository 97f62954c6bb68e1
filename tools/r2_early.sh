# developer: per-kernel times in the contact-poor first steps of the rollout (the driver's --steps 20 --warmup 5 window)
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --kernel-name regex:"Kernel" --launch-skip 100 -c 100 --csv \
  --log-file gpurun_out/r2_early.csv python bench.py --no-cpu --no-steady --steps 20 --warmup 5 > gpurun_out/r2_early.log 2>&1
python - <<'PY'
import csv
from collections import defaultdict
rows=[r for r in csv.reader(open('gpurun_out/r2_early.csv')) if len(r)>5]
h=rows[0]; ki=h.index('Kernel Name'); vi=h.index('Metric Value')
d=defaultdict(list)
for r in rows[1:]:
    try: d[r[ki].split('(')[0]].append(float(r[vi].replace(',',''))/1e3)
    except: pass
tot=0
for k,v in d.items(): print('%-40s n=%d mean=%.1f us min %.1f max %.1f'%(k,len(v),sum(v)/len(v),min(v),max(v))); tot+=sum(v)/len(v)
print('sum', tot)
PY
python bench.py --no-cpu --no-steady --steps 20 --warmup 5 | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('short ms/step', d['ms_per_step'], 'e2e', d['e2e']['value'])"
MYO_CONCURRENT=0 python bench.py --no-cpu --no-steady --steps 20 --warmup 5 | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('short serial ms/step', d['ms_per_step'], 'e2e', d['e2e']['value'])"
