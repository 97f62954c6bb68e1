# developer (GPU box): one full ncu capture (with source counters) of the solver kernel at steady state -> gpurun_out/<tag>_newton.ncu-rep
tag=${1:-r2g}
ncu --set full --import-source on --clock-control none --kernel-name regex:"newtonKernel" --launch-skip 320 -c 1 -f -o gpurun_out/${tag}_newton \
  python bench.py --no-cpu --no-steady --steps 40 --warmup 300 > gpurun_out/${tag}_ncu_newton.log 2>&1
cp myosuite_mjx_b200/csrc/tile.o gpurun_out/${tag}_tile.o
ls -la gpurun_out | grep ${tag}
