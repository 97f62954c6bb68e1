# developer (GPU box): a few memory metrics of every kernel of one steady-state step (warm caches)
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct,lts__t_sectors_srcunit_tex_op_read.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,lts__average_t_sector_hit_rate_srcunit_tex_realtime.pct \
  --clock-control none --cache-control none --kernel-name regex:"Kernel" --launch-skip 4500 -c 15 --csv --log-file gpurun_out/r2_quick_metrics.csv python bench.py --no-cpu --no-steady --steps 40 --warmup 300 > gpurun_out/r2_quick_metrics.log 2>&1
python - <<'PY'
import csv
from collections import OrderedDict
rows=[r for r in csv.reader(open('gpurun_out/r2_quick_metrics.csv')) if len(r)>5]
h=rows[0]; idx={n:i for i,n in enumerate(h)}
out=OrderedDict()
for r in rows[1:]:
    k=r[idx['ID']]; out.setdefault(k,{'name':r[idx['Kernel Name']].split('(')[0].replace('<unnamed>::','')})
    out[k][r[idx['Metric Name']]]=r[idx['Metric Value']]+' '+r[idx['Metric Unit']]
for k,d in out.items():
    if d['name']=='nopKernel': continue
    print('%-22s'%d['name'], ' | '.join('%s=%s'%(m.split('__')[-1][:28],v) for m,v in d.items() if m!='name'))
PY
