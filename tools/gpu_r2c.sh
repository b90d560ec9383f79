#!/bin/bash
# round-2 GPU call C: FP32 pipe-rate microbenchmark, full GPU suite, the no-alloc test alone in a fresh process, bench (both
# arms), ncu launch list + full capture of the fused kernel
cd "$(dirname "$0")/.." || exit 1
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:$LD_LIBRARY_PATH
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/r2c_gpu.txt
timeout 120 tools/ffma_rate.bin > $O/r2c_ffma_rate.jsonl 2>&1
rm -f $O/parity_tests.jsonl $O/closed_loop_tests.jsonl
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/r2c_pytest_noalloc_alone.log 2>&1
echo "pytest rc $?" >> $O/r2c_pytest_noalloc_alone.log
timeout 2400 python -m pytest tests -m gpu -q > $O/r2c_pytest.log 2>&1
echo "pytest rc $?" >> $O/r2c_pytest.log
timeout 900 python bench.py > $O/r2c_bench_n1.json 2> $O/r2c_bench.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > $O/r2c_bench_reference_arm.json 2>> $O/r2c_bench.err
timeout 600 python bench.py --steps 2 --warmup 3 --no-extras > $O/r2c_plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2c_launches.csv python bench.py --steps 2 --warmup 3 --no-extras > $O/r2c_ncu1.log 2>&1
timeout 300 python tools/quick_fused.py --child default > $O/r2c_plain2.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:fused_kernel -s 4 -c 1 -o $O/r2c_fused python tools/quick_fused.py --child default > $O/r2c_ncu2.log 2>&1
echo done
