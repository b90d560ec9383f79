#!/bin/bash
# round-2 GPU call B: operand-split A/B, full GPU test suite, bench (both arms), ncu launch list + fused-kernel capture
cd "$(dirname "$0")/.." || exit 1
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:$LD_LIBRARY_PATH
O=gpurun_out
timeout 600 python tools/quick_fused.py base tmb tmbfd tmbcg1 tmbl8 default > $O/r2b_quick_fused.jsonl 2> $O/r2b_quick_fused.err
rm -f $O/parity_tests.jsonl
timeout 1800 python -m pytest tests -m gpu -q > $O/r2b_pytest.log 2>&1
echo "pytest rc $?" >> $O/r2b_pytest.log
timeout 900 python bench.py > $O/r2b_bench_n1.json 2> $O/r2b_bench.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > $O/r2b_bench_reference_arm.json 2>> $O/r2b_bench.err
timeout 600 python bench.py --steps 2 --warmup 3 --no-extras > $O/r2b_plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2b_launches.csv python bench.py --steps 2 --warmup 3 --no-extras > $O/r2b_ncu1.log 2>&1
timeout 300 python tools/quick_fused.py --child default > $O/r2b_plain2.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:fused_kernel -s 4 -c 1 -o $O/r2b_fused python tools/quick_fused.py --child default > $O/r2b_ncu2.log 2>&1 &&
  timeout 600 ncu --metrics smsp__sass_thread_inst_executed_op_ffma_pred_on.sum,smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,smsp__sass_thread_inst_executed_ops_fadd_fmul_ffma_pred_on.sum,smsp__inst_executed.sum,gpu__time_duration.sum --clock-control none -k regex:fused_kernel -s 4 -c 1 --csv --log-file $O/r2b_fused_flops.csv python tools/quick_fused.py --child default > $O/r2b_ncu3.log 2>&1
echo done
