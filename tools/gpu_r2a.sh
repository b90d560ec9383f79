#!/bin/bash
# round-2 GPU call A: tiling / PCG-variant A/B, operating-point parity report, slot A/B, capped closed loop
cd "$(dirname "$0")/.." || exit 1
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:$LD_LIBRARY_PATH
O=gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/r2a_gpu.txt
timeout 900 python tools/quick_fused.py base k32c2l4 k32c2l8 k64c1l8 k16c4l8 cg1 cg1k32c2l4 > $O/r2a_quick_fused.jsonl 2> $O/r2a_quick_fused.err
rm -f $O/parity_r2.jsonl $O/parity_r2_cg1.jsonl
timeout 900 python tools/parity_report.py --out $O/parity_r2.jsonl --save-fixtures $O > $O/r2a_parity.log 2>&1
MPCGPU_B200_LIB=$PWD/mpcgpu_b200/libexp_cg1.so timeout 900 python tools/parity_report.py --variant 1 --out $O/parity_r2_cg1.jsonl > $O/r2a_parity_cg1.log 2>&1
( cd oracle/_ref/run
  timeout 300 ./slot_ab_n32 5e-6 6 > ../../../$O/r2a_slot_ab.jsonl 2> ../../../$O/r2a_slot_ab.err
  timeout 300 ./slot_ab_n128 1e-4 6 >> ../../../$O/r2a_slot_ab.jsonl 2>> ../../../$O/r2a_slot_ab.err
  timeout 300 ./track_slot_pcg_n32 5e-6 > ../../../$O/r2a_track.jsonl 2> ../../../$O/r2a_track.err
  timeout 300 ./track_slot_b200_n32 5e-6 >> ../../../$O/r2a_track.jsonl 2>> ../../../$O/r2a_track.err
  for cap in 2 4 8; do MPCGPU_B200_SLOT_MAX_SQP=$cap timeout 300 ./track_slot_b200_n32 5e-6 | sed "s/^{/{\"sqp_cap\": $cap, /" >> ../../../$O/r2a_track.jsonl 2>> ../../../$O/r2a_track.err; done
)
echo done
