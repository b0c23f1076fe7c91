#!/bin/bash
# Builds libngp_b200.so (the C-ABI library) for sm_100a.  -fmad=false: see common.cuh.
set -e
cd "$(dirname "$0")"
OUT=../libngp_b200.so
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false -Xcompiler -fPIC -ccbin /usr/bin/g++"
mkdir -p ../build
pids=()
for f in capi_common hash_encode sampler march_train march_infer composite_train render_infer optimizer mlp synthetic_scene l2_probe; do
  if [ ! -f ../build/$f.o ] || [ $f.cu -nt ../build/$f.o ] || [ -n "$(find . -name '*.cuh' -newer ../build/$f.o)" ] || [ ../../include/ngp_b200.h -nt ../build/$f.o ]; then
    $NVCC $FLAGS ${EXTRA:-} -c $f.cu -o ../build/$f.o &
    pids+=($!)
  fi
done
for p in "${pids[@]}"; do wait $p; done
$NVCC -shared -o $OUT ../build/*.o -gencode arch=compute_100a,code=sm_100a -ccbin /usr/bin/g++
echo built $OUT
