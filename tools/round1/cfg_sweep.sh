# developer: rebuild the kernel with different compile-time configurations on the GPU box and time the default bench
cd myosuite_mjx_b200/csrc
for cfg in "-DCTA_WARPS=8 -DCTA_MINBLOCKS=2" "-DCTA_WARPS=8 -DCTA_MINBLOCKS=2 -DNO_PHASE_SYNC" "-DCTA_WARPS=4 -DCTA_MINBLOCKS=4" "-DCTA_WARPS=2 -DCTA_MINBLOCKS=8" "-DCTA_WARPS=1 -DCTA_MINBLOCKS=16 -DNO_PHASE_SYNC"; do
  rm -f team*.o; make -j16 TEAMS="1 2 4 8 16" EXTRA="$cfg" > /dev/null 2>&1
  echo "== $cfg"
  (cd ../..; python bench.py --no-cpu --fields core --steps 150 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('core ms/step %.3f' % d['ms_per_step'], 'env-steps/s %.3e' % d['value'])
")
done
