#!/usr/bin/env python
"""CPU experiment (fp32, numpy + the oracle's block-tridiagonal matvec): can one of the two reductions of a PCG iteration be
hidden behind a matvec by replacing one matvec INPUT with a recurrence, and still meet the operating-point parity criteria
(oracle/parity.py) on the real-IIWA fixtures?

  variant zrec: w = Pinv (S p) is computed while p.Sp is being reduced;  z <- z - alpha w   (instead of z = Pinv r)
  variant srec: S z is computed while r.z is being reduced;              Sp <- S z + beta Sp (instead of Sp = S p)
  optional --replace K: every K iterations the recurrence is replaced by the true matvec.

    python tools/pcg_overlap_experiment.py            (TEST INFRASTRUCTURE: imports oracle/)
"""
import glob
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import parity, bdmv  # noqa: E402

S, C = 14, 7
f32 = np.float32


def run(variant, N, S32, P32, g32, lam0, max_iter, tol, replace=0):
    mv = lambda M, x: bdmv(S, N, M, x.astype(f32)).astype(f32)
    lam = lam0.astype(f32).copy()
    r = (g32 - mv(S32, lam)).astype(f32)
    z = mv(P32, r)
    p = z.copy()
    rz = f32(np.dot(r, z))
    Sp = mv(S32, p)
    it = 0
    if abs(rz) < tol:
        return lam, 0, False
    for it in range(1, max_iter + 1):
        if variant == "srec" and it > 1 and not (replace and it % replace == 0):
            pass                                   # Sp already updated by the recurrence below
        elif it > 1:
            Sp = mv(S32, p)
        if variant == "zrec":
            w = mv(P32, Sp)                        # overlaps the p.Sp reduction
        alpha = f32(rz / f32(np.dot(p, Sp)))
        lam = (lam + alpha * p).astype(f32)
        r = (r - alpha * Sp).astype(f32)
        if variant == "zrec" and not (replace and it % replace == 0):
            z = (z - alpha * w).astype(f32)
        else:
            z = mv(P32, r)
        if variant == "srec":
            Sz = mv(S32, z)                        # overlaps the r.z reduction
        rz_new = f32(np.dot(r, z))
        if abs(rz_new) < tol:
            return lam, it, False
        beta = f32(rz_new / rz)
        rz = rz_new
        p = (z + beta * p).astype(f32)
        if variant == "srec":
            Sp = (Sz + beta * Sp).astype(f32)
    return lam, max_iter, True


def main():
    replace = int(sys.argv[sys.argv.index("--replace") + 1]) if "--replace" in sys.argv else 0
    tol_of = {32: parity.REAL_CASES[0][3], 64: parity.REAL_CASES[2][3], 128: parity.REAL_CASES[4][3]}
    out = {}
    for variant in ("plain", "zrec", "srec"):
        rows = []
        for f in sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "iiwa_kkt_n*_r*.npz"))):
            d = np.load(f)
            N = int(d["N"])
            sysd = parity.systems(N, d["G"], d["C"], d["g"], d["cvec"], float(d["rho"]))
            rng = np.random.default_rng(N + int(d["row0"]))
            warm = (sysd["star"] * (1 + 1e-2 * rng.standard_normal(S * N))).astype(f32)
            for lam0, tag in ((np.zeros(S * N, f32), "cold"), (warm, "warm")):
                for tol in tol_of[N]:
                    lam, it, ex = run(variant, N, sysd["S32"], sysd["P32"], sysd["g32"], lam0, 10000, tol, replace)
                    r = parity.record(N, sysd, lam0, tol, lam, it, ex)
                    r.update(violations=parity.check(r), start=tag, file=os.path.basename(f))
                    rows.append(r)
                    if r["violations"]:
                        print(variant, os.path.basename(f), tag, tol, it, r["iters_oracle_f32"], r["iters_pcg_f64"], r["violations"], flush=True)
        s = parity.summarize(rows)
        out[variant] = {k: s[k] for k in ("sampled", "iters_within_1", "iters_equal_fp64_pcg", "lambda_rel_max", "floor_max", "res_diff_max", "res_floor_max", "violations")}
        print(json.dumps({variant: out[variant], "replace": replace}), flush=True)


if __name__ == "__main__":
    main()
