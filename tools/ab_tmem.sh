#!/bin/bash
# A/B on one box: Pinv rows in tensor memory (default build) vs in shared memory (alternative build
# mpcgpu_b200/libmpcgpu_b200_smem.so = same sources with -DMPCB200_PINV_TMEM=0), 4 vs 8 lanes per knot.
ALT=/root/repo/mpcgpu_b200/libmpcgpu_b200_smem.so
if [ ! -f "$ALT" ]; then   # same sources, Pinv rows kept in shared memory (about 3 minutes of nvcc)
  (cd /root/repo/mpcgpu_b200/csrc && nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC \
     -shared -cudart shared -DMPCB200_PINV_TMEM=0 -o "$ALT" capi.cu capi_lat.cu capi_f64.cu solver.cu) || exit 1
fi
run() { echo "== $1"; shift; env "$@" timeout 300 python bench.py --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);print(json.dumps({'solves_per_s':round(d['value']),'ms_per_step':round(d['ms_per_step'],3),'e2e':round(d['e2e']['value']),'single':[(s['N'],round(s['pcg_us'],1),s['pcg_iters'],round(s['schur_us'],1),round(s['fused_us'],1)) for s in d['single_solve']]}))"; }
run tmem_lpk4
run tmem_lpk8 MPCGPU_B200_LPK=8
run smem_lpk4 MPCGPU_B200_LIB=/root/repo/mpcgpu_b200/libmpcgpu_b200_smem.so
run smem_lpk8 MPCGPU_B200_LIB=/root/repo/mpcgpu_b200/libmpcgpu_b200_smem.so MPCGPU_B200_LPK=8
