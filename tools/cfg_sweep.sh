# developer: rebuild the kernel with different CTA shapes on the GPU box and time them
cd myosuite_mjx_b200/csrc
for cfg in "-DCTA_WARPS=8 -DCTA_MINBLOCKS=2" "-DCTA_WARPS=4 -DCTA_MINBLOCKS=1" "-DCTA_WARPS=4 -DCTA_MINBLOCKS=4" "-DCTA_WARPS=2 -DCTA_MINBLOCKS=1" "-DCTA_WARPS=16 -DCTA_MINBLOCKS=1"; do
  rm -f team*.o; make -j16 EXTRA="$cfg" > /dev/null 2>&1
  echo "== $cfg"
  (cd ../..; for t in 4 8; do MYO_TEAM=$t tools/sweep.sh 8192; done)
done
