#!/bin/bash
# Builds the C++ host layer + pybind11 module pyngp (in-tree, next to libngp_b200.so).  Host code only: g++ against the
# CUDA runtime headers; the kernels are in libngp_b200.so (csrc/build.sh).
set -e
cd "$(dirname "$0")"
PY=${PYTHON:-python}
EXT=$($PY -c "import sysconfig; print(sysconfig.get_config_var('EXT_SUFFIX'))")
PYINC=$($PY -c "import sysconfig; print(sysconfig.get_paths()['include'])")
PBINC=$($PY -c "import pybind11; print(pybind11.get_include())")
NCCL_LIB=$($PY -c "import os, importlib.util as u; s=u.find_spec('nvidia.nccl'); print(os.path.join(list(s.submodule_search_locations)[0],'lib') if s else '')" 2>/dev/null || true)
OUT=../pyngp$EXT
SRCS="python_api.cpp testbed.cpp testbed_train.cpp testbed_render.cpp occupancy_sampler.cpp ngp_network.cpp ray_marcher.cpp dataset.cpp nccl_glue.cpp"
NEWER=0
for f in $SRCS *.h ../../include/ngp_b200.h ../libngp_b200.so; do [ ! -f $OUT ] || [ $f -nt $OUT ] && NEWER=1; done
if [ $NEWER = 1 ]; then
  NCCL_LINK="-lnccl"
  if [ -n "$NCCL_LIB" ] && [ -f "$NCCL_LIB/libnccl.so.2" ]; then NCCL_LINK="-L$NCCL_LIB -l:libnccl.so.2 -Wl,-rpath,$NCCL_LIB"; fi
  mkdir -p ../build/host
  pids=()
  for f in $SRCS; do
    o=../build/host/${f%.cpp}.o
    if [ ! -f $o ] || [ $f -nt $o ] || [ -n "$(find . ../../include -name '*.h' -newer $o)" ]; then
      g++ -O2 -std=c++17 -fPIC -fvisibility=hidden -Wall -Wno-unused-function -I/usr/local/cuda/include -I$PYINC -I$PBINC -c $f -o $o &
      pids+=($!)
    fi
  done
  for p in "${pids[@]}"; do wait $p; done
  objs=""; for f in $SRCS; do objs="$objs ../build/host/${f%.cpp}.o"; done
  g++ -shared -o $OUT $objs -L.. -l:libngp_b200.so -Wl,-rpath,'$ORIGIN' -L/usr/local/cuda/lib64 -lcudart $NCCL_LINK -lz
fi
echo built $OUT
