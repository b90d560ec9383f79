#!/usr/bin/env python
"""Single-solve latency A/B of experiment builds (tools/ab_build.sh): graph-timed PCG / Schur / fused at one N.
   python tools/quick_single.py 128 name1 name2 ..."""
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
        N = int(sys.argv[2])
        # the experiment builds have no reference-style kernel tilings beyond the one under test
        out = bench.single_solve_latency(torch, mg, N, 1e-4)
        print(json.dumps({"build": sys.argv[3], **{k: (round(v, 2) if isinstance(v, float) else v) for k, v in out.items() if k != "timing"}}))
    else:
        N = sys.argv[1]
        for name in sys.argv[2:]:
            lib = os.path.join(ROOT, "mpcgpu_b200", "libexp_%s.so" % name) if name != "default" else \
                os.path.join(ROOT, "mpcgpu_b200", "libmpcgpu_b200.so")
            subprocess.run([sys.executable, os.path.abspath(__file__), "--child", N, name], env=dict(os.environ, MPCGPU_B200_LIB=lib), check=False)
