#!/bin/bash
# multi-GPU bench line (batched weak scaling + e2e + producer e2e + knot-split legs): tools/gpu_multi.sh <ngpus> [tag]
cd "$(dirname "$0")/.." || exit 1
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:$LD_LIBRARY_PATH
N=$1; TAG=${2:-r2}
O=gpurun_out
mkdir -p $O
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 10 --warmup 3 > $O/${TAG}_bench_n$N.json 2> $O/${TAG}_bench_n$N.err
echo "rc $?"; tail -c 600 $O/${TAG}_bench_n$N.err
