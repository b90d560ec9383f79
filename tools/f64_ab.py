"""A/B of the fp64 tilings (MPCGPU_B200_F64_LPK = 8 | 16, MPCGPU_B200_F64_KPC12 = 1 | 0) on batched fused solves."""
import sys, json, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, bench
import mpcgpu_b200 as mg
for N, B in ((64, 1024), (32, 2048), (128, 256)):
    r = bench.f64_leg(torch, mg, N, B, 1e-4)
    print(json.dumps({"lpk": os.environ.get("MPCGPU_B200_F64_LPK", "8"), "kpc12": os.environ.get("MPCGPU_B200_F64_KPC12", "1"), **{k: (round(v, 6) if isinstance(v, float) else v) for k, v in r.items()}}))
