#!/usr/bin/env python
"""Operating-point parity report (VERDICT r1 "Next" #1): the fused GPU solve against the CPU oracle and the
reference's QDLDL on
  * REAL IIWA KKT blocks from the reference's own generate_kkt_submatrices (oracle/_ref/run/ref_harness kkt)
    at N = 32 / 64 / 128, rho = 1e-3, the exit tolerances examples/track_iiwa_pcg.cu:48-68 sweeps, cold and
    warm starts, plus the converged regime (tolerance at the fp32 floor) against QDLDL;
  * instances sampled from the bench batch itself (synth.gen_kkt_fast, 4096 x N=64, tol 1e-4).
One JSON line per solve (oracle/parity.py record + the violated criteria) and one summary line.

    python tools/parity_report.py [--out gpurun_out/parity_r2.jsonl] [--variant 0|1] [--save-fixtures DIR]
GPU box only (needs a GPU and the prebuilt ref_harness); nothing here reads /root/reference.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
RUN = os.path.join(ROOT, "oracle", "_ref", "run")
S, C = 14, 7

def real_kkt(N, row0, perturb, tmp):
    import mpcgpu_b200 as mg
    z = mg.sizes(S, C, N)
    f = os.path.join(tmp, "kkt_%d_%d.bin" % (N, row0))
    env = dict(os.environ, LD_LIBRARY_PATH="/usr/local/cuda/lib64:" + os.environ.get("LD_LIBRARY_PATH", ""))
    subprocess.run([os.path.join(RUN, "ref_harness"), "kkt", str(N), str(row0), str(perturb), f], cwd=RUN, env=env, check=True,
                   capture_output=True, timeout=300)
    buf = np.fromfile(f, np.float32)
    out, o = [], 0
    for n in (z["G"], z["C"], z["g"], z["c"]):
        out.append(buf[o:o + n]); o += n
    return out


def gpu_solve(N, G, Cm, g, cv, rho, lam0, tol, max_iter=10000):
    import torch
    import mpcgpu_b200 as mg
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a, np.float32)).cuda()[None]
    Gd, lam = t(G).clone(), t(lam0).clone()
    it = torch.zeros(1, dtype=torch.int32, device="cuda"); ex = torch.ones(1, dtype=torch.uint8, device="cuda")
    mg.kkt_solve(S, C, N, Gd, t(Cm), t(g), t(cv), rho, lam, it, ex, mg.PcgConfig(pcg_exit_tol=tol, pcg_max_iter=max_iter))
    torch.cuda.synchronize()
    return lam[0].cpu().numpy(), int(it.item()), bool(ex.item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "parity_r2.jsonl"))
    ap.add_argument("--variant", type=int, default=int(os.environ.get("MPCGPU_B200_PCG_VARIANT", "0")))
    ap.add_argument("--save-fixtures", default=None)
    ap.add_argument("--bench-samples", type=int, default=32)
    ap.add_argument("--skip-real", action="store_true")
    a = ap.parse_args()
    from oracle import parity
    from mpcgpu_b200 import synth
    rows = []
    rho = 1e-3
    with tempfile.TemporaryDirectory() as tmp, open(a.out, "a") as out:
        def emit(r):
            out.write(json.dumps(r) + "\n"); out.flush()
        for N, row0, perturb, tols in ([] if a.skip_real else parity.REAL_CASES):
            try:
                G, Cm, g, cv = real_kkt(N, row0, perturb, tmp)
                sysd = parity.systems(N, G, Cm, g, cv, rho)
                rng = np.random.default_rng(N + row0)
                warm = (sysd["star"] * (1 + 1e-2 * rng.standard_normal(S * N))).astype(np.float32)   # SURVEY 8(d), mpcsim.cuh:383
                for lam0, tag in ((np.zeros(S * N, np.float32), "cold"), (warm, "warm")):
                    for tol in tols:
                        lam, it, ex = gpu_solve(N, G, Cm, g, cv, rho, lam0, tol)
                        r = parity.record(N, sysd, lam0, tol, lam, it, ex, a.variant)
                        r.update(source="real_iiwa", row0=row0, perturb=perturb, start=tag, violations=parity.check(r))
                        rows.append(r); emit(r)
                # converged regime: tolerance far below what fp32 can reach, bounded iterations
                lam, it, ex = gpu_solve(N, G, Cm, g, cv, rho, np.zeros(S * N, np.float32), 1e-12, max_iter=2000)
                r = parity.record(N, sysd, np.zeros(S * N, np.float32), 1e-12, lam, it, ex, a.variant, max_iter=2000)
                r.update(source="real_iiwa_converged", row0=row0, perturb=perturb, start="cold", violations=parity.check(r, converged=True))
                emit(r)
                if a.save_fixtures:
                    np.savez_compressed(os.path.join(a.save_fixtures, "iiwa_kkt_n%d_r%d.npz" % (N, row0)), N=N, row0=row0,
                                        perturb=perturb, rho=rho, G=G, C=Cm, g=g, cvec=cv, lam_gpu_tol1e12=lam, iters_gpu_tol1e12=it)
            except Exception as e:      # e.g. an experiment build without this tiling
                emit({"source": "real_iiwa", "N": N, "row0": row0, "error": str(e)[:200]})
        # the bench workload itself
        B, N = 4096, 64
        d = synth.gen_kkt_fast(N, B, seed=synth.BASE_SEED)
        idx = np.linspace(0, B - 1, a.bench_samples).astype(int)
        brow = []
        for b in idx:
            sysd = parity.systems(N, d["G"][b], d["C"][b], d["g"][b], d["c"][b], d["rho"])
            lam0 = np.zeros(S * N, np.float32)
            lam, it, ex = gpu_solve(N, d["G"][b], d["C"][b], d["g"][b], d["c"][b], d["rho"], lam0, 1e-4)
            r = parity.record(N, sysd, lam0, 1e-4, lam, it, ex, a.variant)
            r.update(source="bench_batch", instance=int(b), start="cold", violations=parity.check(r))
            brow.append(r); emit(r)
        emit({"summary_real_iiwa": parity.summarize(rows), "summary_bench_batch": parity.summarize(brow),
              "variant": a.variant, "lib": os.environ.get("MPCGPU_B200_LIB", "default")})
    print(json.dumps({"real": parity.summarize(rows), "bench": parity.summarize(brow)}))


if __name__ == "__main__":
    main()
