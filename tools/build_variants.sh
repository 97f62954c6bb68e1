#!/bin/bash
# developer: build A/B variants of the library into gpurun_variants/lib_<tag>.so (select with MYO_LIB=...); usage: tools/build_variants.sh tag1 "flags1" tag2 "flags2" ...
set -e
cd "$(dirname "$0")/../myosuite_mjx_b200/csrc"
mkdir -p ../../gpurun_variants
NV="/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -prec-div=false -prec-sqrt=false -ftz=true"
make -s -j8 >/dev/null
while [ $# -gt 1 ]; do
  tag=$1; flags=$2; shift 2
  ( $NV $flags -c tile.cu -o /tmp/tile_$tag.o && $NV -shared -o ../../gpurun_variants/lib_$tag.so /tmp/tile_$tag.o plan.o env.o api.o && echo built $tag ) &
done
wait
