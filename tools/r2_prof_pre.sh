# developer (GPU box): full ncu captures (source counters) of the pre-solver kernels of one steady-state step
tag=${1:-r2h}
ncu --set full --import-source on --clock-control none --kernel-name regex:"tendonKernel|inertiaKernel|velactKernel|makeconKernel|collideCompactKernel|posoutKernel|conoutKernel|collideKernel" --launch-skip 2600 -c 8 -f -o gpurun_out/${tag}_pre \
  python bench.py --no-cpu --no-steady --steps 40 --warmup 300 > gpurun_out/${tag}_ncu_pre.log 2>&1
cp myosuite_mjx_b200/csrc/tile.o gpurun_out/${tag}_tile.o
ls -la gpurun_out | grep ${tag}
