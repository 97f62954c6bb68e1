#!/bin/bash
# developer (GPU box): per-kernel launch list + one full capture of every kernel of a steady-state step (phase pipeline)
# usage: tools/r2_prof.sh <tag> [skip] ; 8 launches per step (+1 ctrl draw every 10 steps)
tag=${1:-r2b}
skip=${2:-2440}
out=gpurun_out
python bench.py --no-cpu --steps 20 --warmup 300 > $out/${tag}_plain.json 2> $out/${tag}_plain.err || { echo "plain run failed"; tail -5 $out/${tag}_plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.per_cycle_active,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum \
  --clock-control none --cache-control none --launch-skip $skip -c 20 --csv --log-file $out/${tag}_launches.csv python bench.py --no-cpu --steps 20 --warmup 300 > $out/${tag}_ncu1.log 2>&1
if [ "$3" == "full" ]; then
ncu --set full --import-source on --clock-control none --launch-skip $skip -c 9 -f -o $out/${tag}_full python bench.py --no-cpu --steps 20 --warmup 300 > $out/${tag}_ncu2.log 2>&1
fi
python - <<PY
import csv
rows=list(csv.reader(open("$out/${tag}_launches.csv")))
h=next(i for i,r in enumerate(rows) if r and r[0]=="ID")
H=rows[h]; ik=H.index("Kernel Name"); im=H.index("Metric Name"); iv=H.index("Metric Value"); iid=H.index("ID")
d={}
for r in rows[h+1:]:
    d.setdefault((int(r[iid]), r[ik].split("(")[0][-28:]), {})[r[im]]=float(r[iv].replace(",",""))
tot=0
for (i,k),m in sorted(d.items()):
    t=m.get("gpu__time_duration.sum",0)/1e3; tot+=t
    print(f"{i:3d} {k:28s} {t:8.1f} us  rd {m.get('dram__bytes_read.sum',0)/1e6:7.1f} MB wr {m.get('dram__bytes_write.sum',0)/1e6:7.1f} MB  issue {m.get('smsp__issue_active.avg.pct_of_peak_sustained_active',0):5.1f}%  warps {m.get('sm__warps_active.avg.per_cycle_active',0):5.1f}  thr/inst {m.get('smsp__thread_inst_executed_per_inst_executed.ratio',0):5.1f}  inst {m.get('smsp__inst_executed.sum',0)/1e6:7.2f} M")
print("total us", tot)
PY
