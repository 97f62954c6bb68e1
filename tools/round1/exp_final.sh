#!/bin/bash
# developer (GPU box): env-layer throughput with the final kernels + per-function profile of the two-CTA build at 16384 worlds
python tools/env_bench.py 8192 100 > gpurun_out/r1zz_env_bench.jsonl 2> gpurun_out/r1zz_env_bench.err
python tools/env_bench.py 16384 100 > gpurun_out/r1zz_env_bench_16384.jsonl 2>> gpurun_out/r1zz_env_bench.err
cp myosuite_mjx_b200/csrc/team8x2.o gpurun_out/r1zz_team8x2.o
timeout 600 ncu --set full --import-source on --clock-control none -k regex:stepKernel --launch-skip 300 -c 1 -f -o gpurun_out/r1zz_x2 \
  python bench.py --no-cpu --fields core --nworld 16384 --steps 400 --warmup 100 > gpurun_out/r1zz_x2.log 2>&1
cat gpurun_out/r1zz_env_bench.jsonl | cut -c1-200
