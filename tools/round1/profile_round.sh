#!/bin/bash
# developer (GPU box): the round's evidence set -> gpurun_out/<tag>_*   usage: tools/profile_round.sh r1x
tag=${1:-r1x}
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_bench_reference.json 2> gpurun_out/${tag}_ref.err
python bench.py > gpurun_out/${tag}_bench_n1.json 2> gpurun_out/${tag}_bench.err
python bench.py --no-cpu --fields core > gpurun_out/${tag}_bench_core.json 2>> gpurun_out/${tag}_bench.err
python bench.py --no-cpu --nworld 16384 > gpurun_out/${tag}_bench_n1_16384.json 2>> gpurun_out/${tag}_bench.err
# launch list (steady state: the last 20 steps of a 500-step rollout), only after the plain run above exited 0
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --launch-skip 960 -c 40 --csv \
  --log-file gpurun_out/${tag}_launches.csv python bench.py --no-cpu --steps 400 --warmup 100 > gpurun_out/${tag}_ncu1.log 2>&1
# one full capture of each kernel at steady state
ncu --set full --import-source on --clock-control none --launch-skip 800 -c 2 -f -o gpurun_out/${tag}_full \
  python bench.py --no-cpu --steps 400 --warmup 100 > gpurun_out/${tag}_ncu2.log 2>&1
cp myosuite_mjx_b200/csrc/team8.o gpurun_out/${tag}_team8.o
ls -la gpurun_out/ | grep ${tag}
