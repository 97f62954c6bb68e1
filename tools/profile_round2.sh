#!/bin/bash
# developer (GPU box): the round-2 evidence set -> gpurun_out/<tag>_*   usage: tools/profile_round2.sh r2z
tag=${1:-r2z}
out=gpurun_out
python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_reference.json 2> $out/${tag}_ref.err
python bench.py > $out/${tag}_bench_n1.json 2> $out/${tag}_bench.err
python bench.py --steps 20 --warmup 5 --no-cpu > $out/${tag}_bench_short.json 2>> $out/${tag}_bench.err
python bench.py --no-cpu --fields core > $out/${tag}_bench_core.json 2>> $out/${tag}_bench.err
python bench.py --no-cpu --nworld 16384 > $out/${tag}_bench_n1_16384.json 2>> $out/${tag}_bench.err
python bench.py --no-cpu --nworld 32768 --steps 400 > $out/${tag}_bench_n1_32768.json 2>> $out/${tag}_bench.err
python bench.py --no-cpu --model legs > $out/${tag}_bench_legs.json 2>> $out/${tag}_bench.err
python bench.py --no-cpu --model hand_sar --steps 300 --no-steady > $out/${tag}_bench_hand_sar.json 2>> $out/${tag}_bench.err
python bench.py --no-cpu --model finger --nworld 4096 > $out/${tag}_bench_finger4096.json 2>> $out/${tag}_bench.err
python bench.py --no-cpu --model elbow --nworld 65536 --steps 300 > $out/${tag}_bench_elbow65536.json 2>> $out/${tag}_bench.err
# launch list at steady state (14 launches per step incl. 2 empty ordering launches; the ctrl draw every 10 steps is torch's), the recipe's pass: cold caches, serialised
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.per_cycle_active,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum \
  --clock-control none --kernel-name regex:"Kernel" --launch-skip 4500 -c 60 --csv --log-file $out/${tag}_launches.csv python bench.py --no-cpu --no-steady --steps 40 --warmup 300 > $out/${tag}_ncu1.log 2>&1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum \
  --clock-control none --cache-control none --kernel-name regex:"Kernel" --launch-skip 4500 -c 60 --csv --log-file $out/${tag}_launches_warm.csv python bench.py --no-cpu --no-steady --steps 40 --warmup 300 > $out/${tag}_ncu1w.log 2>&1
# one full capture of every kernel of a steady-state step
ncu --set full --import-source on --clock-control none --kernel-name regex:"Kernel" --launch-skip 4500 -c 15 -f -o $out/${tag}_full \
  python bench.py --no-cpu --no-steady --steps 40 --warmup 300 > $out/${tag}_ncu2.log 2>&1
cp myosuite_mjx_b200/csrc/tile.o $out/${tag}_tile.o
ls -la $out/ | grep ${tag}
for f in n1 short core n1_16384 n1_32768 legs hand_sar finger4096 elbow65536; do python - <<PY
import json
try:
    d = json.loads(open("$out/${tag}_bench_$f.json").read().strip().splitlines()[-1])
    ss = d.get("steady_state") or {}
    print("$f: ms/step %.3f value %.3e e2e %.3e steady %.3e frac %.4f steady_frac %s flagged %d" % (d["ms_per_step"], d["value"], d["e2e"]["value"], ss.get("value", 0), d["roofline"]["frac"], d["roofline"].get("steady_state_frac"), d["config"]["worlds_with_status_flags"]))
except Exception as e:
    print("$f: FAILED", e)
PY
done
tail -3 $out/${tag}_bench.err
