#!/usr/bin/env python
"""A/B of library builds on the bench workload (4096 x N=64 fused solves, resident inputs):
   python tools/quick_fused.py base wpc2 ...   -> one JSON line per mpcgpu_b200/libexp_<name>.so
Each build runs in its own process (MPCGPU_B200_LIB).  `lam_sha`/`iters_sum` show whether two builds
produce the same bits."""
import hashlib
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def child(name):
    sys.path.insert(0, ROOT)
    import numpy as np
    import torch
    import mpcgpu_b200 as mg
    from mpcgpu_b200 import synth
    S, C, N, B = 14, 7, int(os.environ.get("QF_N", "64")), int(os.environ.get("QF_BATCH", "4096"))
    d = synth.gen_kkt_fast(N, B, seed=synth.BASE_SEED)
    dev = "cuda"
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    G0, Cm, g, cv = t(d["G"]), t(d["C"]), t(d["g"]), t(d["c"])
    G = G0.clone()
    lam = torch.zeros((B, S * N), device=dev)
    it = torch.zeros(B, dtype=torch.int32, device=dev); ex = torch.ones(B, dtype=torch.uint8, device=dev)
    cfg = mg.PcgConfig(pcg_exit_tol=float(os.environ.get("QF_TOL", "1e-4")), pcg_max_iter=int(os.environ.get("QF_MAXIT", "10000")))
    ts = []
    for rep in range(13):
        G.copy_(G0); lam.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        mg.kkt_solve(S, C, N, G, Cm, g, cv, d["rho"], lam, it, ex, cfg, batch=B)
        e1.record(); torch.cuda.synchronize()
        if rep >= 3:
            ts.append(e0.elapsed_time(e1))
    ms = float(np.mean(ts))
    print(json.dumps({"build": name, "kernel_ms": round(ms, 4), "min_ms": round(min(ts), 4), "solves_per_s": round(B / ms * 1e3),
                      "iters_sum": int(it.sum().item()), "exits": int(ex.sum().item()),
                      "lam_sha": hashlib.sha1(lam.cpu().numpy().tobytes()).hexdigest()[:12],
                      "ginv_sha": hashlib.sha1(G.cpu().numpy().tobytes()).hexdigest()[:12]}))


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--child":
        child(sys.argv[2])
    else:
        for name in sys.argv[1:]:
            lib = os.path.join(ROOT, "mpcgpu_b200", "libexp_%s.so" % name) if name != "default" else \
                os.path.join(ROOT, "mpcgpu_b200", "libmpcgpu_b200.so")
            env = dict(os.environ, MPCGPU_B200_LIB=lib)
            subprocess.run([sys.executable, os.path.abspath(__file__), "--child", name], env=env, check=False)
