#!/usr/bin/env python
"""How far is a library build's PCG iterate from the exact-arithmetic iterate (fp64 run of the reference's iteration, same
system, same iteration count), next to the CPU oracle's fp32 runs of the same iterations?  One JSON line per (N, tol) case and
build; used to compare the default library (PCG variant 2 on the single-CTA tilings) with a variant-0 build.
    python tools/variant_drift.py default fullv0      (GPU box; TEST INFRASTRUCTURE: imports oracle/)"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASES = [(8, 1e-6), (16, 1e-6), (16, 1e-5), (32, 1e-6), (33, 1e-6), (64, 1e-6), (64, 1e-5), (64, 1e-4), (32, 1e-4)]
BATCH = int(os.environ.get("VD_BATCH", "160"))      # copies of the system in one call: > 148 CTAs selects the single-CTA (throughput) tilings


def child(name):
    sys.path.insert(0, ROOT)
    import numpy as np
    import torch
    import mpcgpu_b200 as mg
    import oracle
    from mpcgpu_b200 import synth
    S, C = 14, 7
    rel = lambda a, b: float(np.linalg.norm(np.asarray(a, np.float64) - np.asarray(b, np.float64)) / np.linalg.norm(np.asarray(b, np.float64)))
    for seed_off in (0, 1):
        for N, tol in CASES:
            d = synth.gen_kkt(N, 1, seed=synth.BASE_SEED + 3 * N + 1000 * seed_off)
            S32, P32, g32, _ = oracle.schur(S, C, N, d["G"][0], d["C"][0], d["g"][0], d["c"][0], d["rho"], np.float32)
            t = lambda a: torch.from_numpy(np.ascontiguousarray(a, np.float32)).cuda()[None].repeat(BATCH, 1).contiguous()
            lam = torch.zeros((BATCH, S * N), device="cuda")
            it = torch.zeros(BATCH, dtype=torch.int32, device="cuda"); ex = torch.ones(BATCH, dtype=torch.uint8, device="cuda")
            mg.pcg(S, N, t(S32), t(P32), t(g32), lam, it, ex, mg.PcgConfig(pcg_exit_tol=tol, pcg_max_iter=10000), batch=BATCH)
            torch.cuda.synchronize()
            assert bool((lam == lam[0]).all()) and bool((it == it[0]).all()), "copies of one system must give the same bits"
            k = int(it[0].item())
            lam_g = lam[0].cpu().numpy()
            f64 = lambda a: np.asarray(a, np.float64)
            l64, _, _ = oracle.pcg(S, N, f64(S32), f64(P32), f64(g32), np.zeros(S * N), k, 0.0, np.float64)
            l32, _, _ = oracle.pcg(S, N, S32, P32, g32, np.zeros(S * N, np.float32), k, 0.0)
            l32b, _, _ = oracle.pcg(S, N, S32, P32, g32, np.zeros(S * N, np.float32), k, 0.0, np.float32, mode=1)
            lsp, _, _ = oracle.pcg_sprec(S, N, S32, P32, g32, np.zeros(S * N, np.float32), k, 0.0)
            _, k64, _ = oracle.pcg(S, N, f64(S32), f64(P32), f64(g32), np.zeros(S * N), 10000, tol, np.float64)
            res = lambda l: float(np.linalg.norm(oracle.bdmv(S, N, f64(S32), f64(l)) - g32) / np.linalg.norm(g32))
            print(json.dumps({"build": name, "batch": BATCH, "lpk": os.environ.get("MPCGPU_B200_LPK", "4"), "N": N, "tol": tol, "seed": seed_off, "iters_gpu": k, "iters_f64": int(k64),
                              "gpu_vs_f64": rel(lam_g, l64), "oracle_f32_vs_f64": rel(l32, l64), "oracle_f32alt_vs_f64": rel(l32b, l64),
                              "oracle_sprec_vs_f64": rel(lsp, l64), "res_gpu": res(lam_g), "res_f64": res(l64)}), flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--child":
        child(sys.argv[2])
    else:
        for name in sys.argv[1:]:
            lib = os.path.join(ROOT, "mpcgpu_b200", "libexp_%s.so" % name) if name != "default" else \
                os.path.join(ROOT, "mpcgpu_b200", "libmpcgpu_b200.so")
            subprocess.run([sys.executable, os.path.abspath(__file__), "--child", name], env=dict(os.environ, MPCGPU_B200_LIB=lib), check=False)
