import sys, json, os
sys.path.insert(0, '/root/repo')
import torch, bench
import mpcgpu_b200 as mg
for N, B in ((64, 1024), (32, 2048), (128, 256)):
    r = bench.f64_leg(torch, mg, N, B, 1e-4)
    print(json.dumps({"lpk": os.environ.get("MPCGPU_B200_F64_LPK", "16"), **{k: (round(v, 3) if isinstance(v, float) else v) for k, v in r.items()}}))
