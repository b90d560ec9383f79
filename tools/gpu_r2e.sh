#!/bin/bash
# round-2 GPU call E: A/B of the final kernel variants, full GPU suite (new: L1 stage-1 test, determinism stress tests),
# bench, and the ncu launch list of the reference's closed loop (N = 128) on the full L1 shim
cd "$(dirname "$0")/.." || exit 1
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:$LD_LIBRARY_PATH
O=gpurun_out
mkdir -p $O
timeout 600 python tools/quick_fused.py default cg1 fastdiv l8 k32c2 > $O/r2e_quick_fused.jsonl 2> $O/r2e_quick_fused.err
rm -f $O/parity_tests.jsonl $O/closed_loop_tests.jsonl
timeout 2700 python -m pytest tests -m gpu -q > $O/r2e_pytest.log 2>&1
echo "pytest rc $?" >> $O/r2e_pytest.log
timeout 900 python bench.py > $O/r2e_bench_n1.json 2> $O/r2e_bench.err
( cd oracle/_ref/run
  timeout 300 ./track_slot_pcg_n128 1e-4 > ../../../$O/r2e_l1_plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 2000 -c 60 --csv --log-file ../../../$O/r2e_l1_track_n128_launches.csv ./track_slot_pcg_n128 1e-4 > ../../../$O/r2e_l1_ncu.log 2>&1 )
echo done
