#!/bin/bash
# round-2 GPU call F: A/B of the L2 input prefetch and of PCG variant 2 (S p by recurrence, r.rt reduction hidden behind
# the matvec) on the bench workload; operating-point parity report of the variant-2 build
cd "$(dirname "$0")/.." || exit 1
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:$LD_LIBRARY_PATH
O=gpurun_out
mkdir -p $O
timeout 600 python tools/quick_fused.py default pf v2 "$@" > $O/r2f_quick_fused.jsonl 2> $O/r2f_quick_fused.err
MPCGPU_B200_L2_PREFETCH=0 timeout 300 python tools/quick_fused.py pf v2 >> $O/r2f_quick_fused.jsonl 2>> $O/r2f_quick_fused.err
QF_MAXIT=80 QF_TOL=0 timeout 300 python tools/quick_fused.py default v2 >> $O/r2f_quick_fused.jsonl 2>> $O/r2f_quick_fused.err
rm -f $O/r2f_parity_v2.jsonl
MPCGPU_B200_LIB=$PWD/mpcgpu_b200/libexp_v2.so timeout 900 python tools/parity_report.py --out $O/r2f_parity_v2.jsonl > $O/r2f_parity_v2.summary 2> $O/r2f_parity_v2.err
echo done
