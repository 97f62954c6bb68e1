#!/bin/bash
# developer (GPU box): per-function ncu profile of the 28-warps-per-SM build (T = 16, two co-resident 14-warp CTAs, <= 72 registers)
(cd myosuite_mjx_b200/csrc; rm -f team16.o; make EXTRA="-DCTA_WARPS=14 -DCTA_MINBLOCKS=2" > /dev/null 2>&1)
cp myosuite_mjx_b200/csrc/team16.o gpurun_out/team16_14x2.o
MYO_TEAM=16 timeout 600 ncu --set full --import-source on --clock-control none -k regex:stepKernel --launch-skip 300 -c 1 -f -o gpurun_out/prof72 \
  python bench.py --no-cpu --fields core --steps 400 --warmup 100 > gpurun_out/prof72.log 2>&1
(cd myosuite_mjx_b200/csrc; rm -f team16.o; make > /dev/null 2>&1)
ls -la gpurun_out/prof72.ncu-rep
