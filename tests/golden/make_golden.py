"""Generates the committed golden fixtures (run HERE, where /root/reference exists).

1. qdldl_kats.json -- the 9 known-answer tests of the reference's vendored QDLDL
   (qdldl/tests/test_*.h: Ap, Ai, Ax, b, xsol or "must fail"), parsed from the headers.
   They pin oracle/_ref (the reference's own solver build) on the GPU box.
2. kkt_n4.npz / kkt_n8_warm.npz -- small KKT instances: inputs, the Schur system from the
   oracle restatement in fp64, and lambda* from the REFERENCE QDLDL (double build) on it.
   Pins the oracle against drift and gives the GPU tests a reference-derived vector.

Usage:  python tests/golden/make_golden.py
"""
import json
import os
import re
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF_TESTS = "/root/reference/qdldl/tests"


def _arr(src, name):
    m = re.search(r"%s\s*\[\s*\]\s*=\s*\{([^}]*)\}" % name, src)
    if not m:
        return None
    return [float(x) for x in m.group(1).replace("\n", " ").split(",") if x.strip()]


def parse_kats():
    kats = []
    for fn in sorted(os.listdir(REF_TESTS)):
        if not (fn.startswith("test_") and fn.endswith(".h")):
            continue
        src = open(os.path.join(REF_TESTS, fn)).read()
        src = re.sub(r"//[^\n]*", "", src)
        kat = {
            "name": fn[:-2],
            "source": "qdldl/tests/" + fn,
            "Ap": [int(v) for v in _arr(src, "Ap")],
            "Ai": [int(v) for v in _arr(src, "Ai")],
            "Ax": _arr(src, "Ax"),
            "b": _arr(src, "b"),
            "xsol": _arr(src, "xsol"),
            "must_fail": "status < 0" in src,
            "tol": 1e-4,  # QDLDL_TESTS_TOL, qdldl/tests/minunit.h:13
        }
        kats.append(kat)
    return kats


def make_kkt(name, N, seed, warm):
    import oracle
    from mpcgpu_b200 import synth
    s, c = 14, 7
    d = synth.gen_kkt(N, 1, seed=seed)
    G, C, g, cv = d["G"][0], d["C"][0], d["g"][0], d["c"][0]
    S64, P64, gam64, Gi64 = oracle.schur(s, c, N, G, C, g, cv, d["rho"], np.float64)
    lam_star = oracle.qdldl_solve(s, N, S64, gam64)            # reference QDLDL, double
    S32, P32, gam32, Gi32 = oracle.schur(s, c, N, G, C, g, cv, d["rho"], np.float32)
    lam_q32 = oracle.qdldl_solve(s, N, S32, gam32)             # reference QDLDL, float (its own config)
    lam0 = np.zeros(s * N, np.float32)
    if warm:
        rng = np.random.default_rng(seed)
        lam0 = (lam_star * (1 + 1e-2 * rng.standard_normal(s * N))).astype(np.float32)
    lam_pcg, iters, flag = oracle.pcg(s, N, S32, P32, gam32, lam0, 10000, 1e-6, np.float32)
    np.savez_compressed(
        os.path.join(HERE, name), N=N, s=s, c=c, rho=d["rho"], G=G, C=C, g=g, cvec=cv,
        S64=S64, Pinv64=P64, gamma64=gam64, Ginv64=Gi64, lam_star=lam_star, lam_qdldl_f32=lam_q32,
        lam0=lam0, pcg_tol=1e-6, pcg_lambda_f32=lam_pcg, pcg_iters=iters, pcg_exit=flag)


if __name__ == "__main__":
    kats = parse_kats()
    assert len(kats) == 9, len(kats)
    json.dump(kats, open(os.path.join(HERE, "qdldl_kats.json"), "w"), indent=1)
    make_kkt("kkt_n4.npz", 4, 0x4D504347, warm=False)
    make_kkt("kkt_n8_warm.npz", 8, 0x4D504347 + 7, warm=True)
    print("wrote", os.listdir(HERE))
