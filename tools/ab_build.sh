#!/bin/bash
# Experiment build: mpcgpu_b200/libexp_<name>.so = the library restricted to the bench tiling
# (-DMPCB200_MINIMAL: Cfg<64,1,4> only) plus the given nvcc flags.  About a minute of nvcc.
#   tools/ab_build.sh wpc2 -DMPCB200_TM_WPC=2
name=$1; shift
cd "$(dirname "$0")/../mpcgpu_b200/csrc" || exit 1
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -shared -cudart shared \
  -DMPCB200_MINIMAL "$@" -Xptxas -v -o ../libexp_$name.so capi.cu solver.cu 2>&1 | grep -A2 "fused_kernel" | grep -E "registers|spill" 
ls -la ../libexp_$name.so
