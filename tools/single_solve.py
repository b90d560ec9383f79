"""Single-instance solves (latency tiling) for profiling: python tools/single_solve.py [N] [reps]"""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mpcgpu_b200 as mg
from mpcgpu_b200 import synth

N = int(sys.argv[1]) if len(sys.argv) > 1 else 128
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
S, C = 14, 7
d = synth.gen_kkt(N, 1, seed=synth.BASE_SEED + N)
dev = "cuda"
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
G0, Cm, g, cv = t(d["G"]), t(d["C"]), t(d["g"]), t(d["c"])
z = mg.sizes(S, C, N)
Sg = torch.empty((1, z["S"]), device=dev); Pg = torch.empty_like(Sg); gam = torch.empty((1, S * N), device=dev)
G = G0.clone()
mg.form_schur_system(S, C, N, G, Cm, g, cv, Sg, Pg, gam, d["rho"])
lam = torch.zeros((1, S * N), device=dev)
it = torch.zeros(1, dtype=torch.int32, device=dev); ex = torch.zeros(1, dtype=torch.uint8, device=dev)
cfg = mg.PcgConfig(pcg_exit_tol=1e-4)
for _ in range(reps):
    lam.zero_()
    mg.pcg(S, N, Sg, Pg, gam, lam, it, ex, cfg)
torch.cuda.synchronize()
# GPU-bound timing: a CUDA graph of `reps` (reset + solve) pairs
st = torch.cuda.Stream()
with torch.cuda.stream(st):
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr, stream=st):
        for _ in range(20):
            lam.zero_()
            mg.pcg(S, N, Sg, Pg, gam, lam, it, ex, cfg)
    gr.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st); gr.replay(); e1.record(st); torch.cuda.synchronize()
t_full = e0.elapsed_time(e1) / 20 * 1e3
iters = int(it.item())
cfg0 = mg.PcgConfig(pcg_exit_tol=1e-4, pcg_max_iter=0)
with torch.cuda.stream(st):
    g0 = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g0, stream=st):
        for _ in range(20):
            lam.zero_()
            mg.pcg(S, N, Sg, Pg, gam, lam, it, ex, cfg0)
    g0.replay(); torch.cuda.synchronize()
    e0.record(st); g0.replay(); e1.record(st); torch.cuda.synchronize()
t0 = e0.elapsed_time(e1) / 20 * 1e3
print("N %d iters %d  solve %.1f us  floor(max_iter=0) %.1f us  per-iter %.3f us" % (N, iters, t_full, t0, (t_full - t0) / max(iters, 1)))
