#!/bin/bash
# round-2 GPU call G: ncu --set full capture (with source) of one launch of an experiment build of the fused kernel
#   bash tools/gpu_r2g.sh <build name> [tag]
cd "$(dirname "$0")/.." || exit 1
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:$LD_LIBRARY_PATH
B=${1:-v2}; TAG=${2:-r2g}
O=gpurun_out
mkdir -p $O
LIB=$PWD/mpcgpu_b200/libexp_$B.so
[ "$B" = default ] && LIB=$PWD/mpcgpu_b200/libmpcgpu_b200.so
MPCGPU_B200_LIB=$LIB timeout 300 python tools/quick_fused.py --child $B > $O/${TAG}_plain.jsonl 2> $O/${TAG}_plain.err &&
MPCGPU_B200_LIB=$LIB timeout 900 ncu --set full --clock-control none --import-source on -k regex:fused_kernel -s 4 -c 1 -f -o $O/${TAG}_fused python tools/quick_fused.py --child $B > $O/${TAG}_ncu.log 2>&1
echo done
