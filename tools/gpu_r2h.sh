#!/bin/bash
# round-2 GPU call H: full GPU suite on the final default (PCG variant 0; 16 B vector loads, radix
# warp sums, L2 input prefetch), parity report, bench (+ reference arm), ncu launch lists (bench; the reference's closed loop
# on the L1 shim with the SQP wall-clock budget lifted) and one full capture of the fused kernel
cd "$(dirname "$0")/.." || exit 1
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:$LD_LIBRARY_PATH
O=gpurun_out; T=${1:-r2h}
mkdir -p $O
rm -f $O/parity_tests.jsonl $O/closed_loop_tests.jsonl $O/${T}_parity.jsonl
timeout 2700 python -m pytest tests -m gpu -q > $O/${T}_pytest.log 2>&1
echo "pytest rc $?" >> $O/${T}_pytest.log
timeout 900 python tools/parity_report.py --out $O/${T}_parity.jsonl > $O/${T}_parity.summary 2> $O/${T}_parity.err
timeout 900 python bench.py > $O/${T}_bench_n1.json 2> $O/${T}_bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/${T}_bench_reference.json 2>> $O/${T}_bench.err
timeout 300 python tools/quick_fused.py default > $O/${T}_quick_fused.jsonl 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${T}_launches.csv python bench.py --steps 2 --warmup 3 --no-extras > $O/${T}_ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fused_kernel -s 4 -c 1 -f -o $O/${T}_fused python tools/quick_fused.py --child default > $O/${T}_ncu2.log 2>&1
bash tools/gpu_r2i.sh ${T} > $O/${T}_r2i.log 2>&1
echo done
