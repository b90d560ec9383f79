#!/usr/bin/env python
"""Long horizons on one GPU (knot-split kernel, all ranges resident), A/B of library builds:
   python tools/quick_long.py default flat   -> one JSON line per build and horizon (bench.long_horizon_one_gpu)"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

if __name__ == "__main__":
    if sys.argv[1] == "--child":
        sys.path.insert(0, ROOT)
        import torch
        import mpcgpu_b200 as mg
        import bench
        for n, c in ((512, 2), (1024, 2), (2048, 4)):
            out = bench.long_horizon_one_gpu(torch, mg, torch.device("cuda:0"), n, c, 1e-4)
            print(json.dumps({"build": sys.argv[2], **{k: (round(v, 3) if isinstance(v, float) else v) for k, v in out.items()}}), flush=True)
    else:
        for name in sys.argv[1:]:
            lib = os.path.join(ROOT, "mpcgpu_b200", "libexp_%s.so" % name) if name != "default" else \
                os.path.join(ROOT, "mpcgpu_b200", "libmpcgpu_b200.so")
            subprocess.run([sys.executable, os.path.abspath(__file__), "--child", name], env=dict(os.environ, MPCGPU_B200_LIB=lib), check=False)
