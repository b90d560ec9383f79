#!/usr/bin/env python
"""Small, bounded workload for compute-sanitizer (racecheck / synccheck / memcheck; ONE tool per gpurun call, see
B200_PROFILING.md): every kernel family of the library once, few PCG iterations, every cluster size.
    compute-sanitizer --tool racecheck python tools/sanitize_target.py
Cases: fused + two-stage + dz at CL = 1 (throughput tiling, 64 knots per CTA), CL = 4 / 8 / 16 (latency tilings),
a ragged horizon (N = 100), the fp64 twins, and the knot-range split emulated on one GPU (world 4, cluster of 2)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch            # noqa: E402
import mpcgpu_b200 as mg    # noqa: E402
from mpcgpu_b200 import synth   # noqa: E402

S, C = 14, 7
DEV = "cuda:0"
MAX_ITER = int(os.environ.get("SAN_MAX_ITER", "12"))


def t(a, dt=None):
    x = torch.from_numpy(np.ascontiguousarray(a))
    return (x.to(dt) if dt else x).to(DEV)


def run_case(N, B, latency, dtype=torch.float32):
    os.environ["MPCGPU_B200_LATENCY_TILING"] = "1" if latency else "0"
    d = synth.gen_kkt_fast(N, B, seed=7 + N)
    z = mg.sizes(S, C, N)
    G, Cm, g, cv = t(d["G"], dtype), t(d["C"], dtype), t(d["g"], dtype), t(d["c"], dtype)
    G0 = G.clone()
    Sg = torch.empty((B, z["S"]), device=DEV, dtype=dtype); Pg = torch.empty_like(Sg)
    gam = torch.empty((B, z["gamma"]), device=DEV, dtype=dtype)
    lam = torch.zeros((B, S * N), device=DEV, dtype=dtype)
    dz = torch.empty((B, z["g"]), device=DEV, dtype=dtype)
    it = torch.zeros(B, dtype=torch.int32, device=DEV); ex = torch.zeros(B, dtype=torch.uint8, device=DEV)
    cfg = mg.PcgConfig(pcg_exit_tol=1e-30, pcg_max_iter=MAX_ITER)
    mg.form_schur_system(S, C, N, G, Cm, g, cv, Sg, Pg, gam, d["rho"], batch=B)
    mg.pcg(S, N, Sg, Pg, gam, lam, it, ex, cfg, batch=B)
    mg.compute_dz(S, C, N, G, Cm, g, lam, dz, batch=B)
    G.copy_(G0); lam.zero_()
    mg.kkt_solve_dz(S, C, N, G, Cm, g, cv, d["rho"], lam, dz, it, ex, cfg, batch=B)
    torch.cuda.synchronize()
    assert bool(torch.isfinite(lam).all()) and int(it.min()) == MAX_ITER
    return "N=%d B=%d latency=%d %s ok" % (N, B, latency, str(dtype).split(".")[-1])


def run_split(N, world):
    import oracle
    d = synth.gen_kkt(N, 1, seed=5)
    S32, P32, g32, _ = oracle.schur(S, C, N, d["G"][0], d["C"][0], d["g"][0], d["c"][0], d["rho"], np.float32)
    mail = torch.zeros((world, mg.mailbox_bytes() // 4), dtype=torch.int32, device=DEV)
    ptrs = [mail[r].data_ptr() for r in range(world)]
    it = torch.zeros(world, dtype=torch.int32, device=DEV); ex = torch.zeros(world, dtype=torch.uint8, device=DEV)
    cfg = mg.PcgConfig(pcg_exit_tol=1e-30, pcg_max_iter=MAX_ITER)
    epoch = 0
    for _ in range(2):
        lam = torch.zeros(S * N, device=DEV)
        mg.knotsplit_pcg(S, N // world, world, 0, world, epoch, ptrs, t(S32), t(P32), t(g32), lam, it, ex, cfg)
        epoch += mg.epoch_stride(cfg)
    torch.cuda.synchronize()
    assert bool(torch.isfinite(lam).all())
    return "knot split N=%d world=%d ok" % (N, world)


if __name__ == "__main__":
    which = sys.argv[1:] or ["all"]
    out = []
    if "all" in which or "fused" in which:
        out.append(run_case(64, 3, latency=False))            # CL = 1, 64 knots per CTA, TMEM
        out.append(run_case(64, 2, latency=True))             # CL = 4
        out.append(run_case(100, 2, latency=False))           # ragged, CL = 2
        out.append(run_case(128, 1, latency=True))            # CL = 8
        out.append(run_case(256, 1, latency=True))            # CL = 16
        out.append(run_case(32, 2, latency=False))
    if "all" in which or "f64" in which:
        out.append(run_case(64, 2, latency=True, dtype=torch.float64))
    if "all" in which or "split" in which:
        out.append(run_split(128, 4))
    print("\n".join(out))
