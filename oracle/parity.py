"""TEST INFRASTRUCTURE ONLY -- the operating-point parity record of one KKT solve.

north_star: "Results must match the reference's own PCG and QDLDL solves on the same KKT inputs:
iteration count within +-1, and solution and residual within 1e-4 relative".  What that means at the
reference's OWN operating point (rho = 1e-3, include/mpcsim.cuh:256; exit tolerances of
examples/track_iiwa_pcg.cu:48-68) is decided by two facts this module measures for every input:

  * the PCG there is a TRUNCATED solve: an fp64 run of the same iteration stops at the same count with
    the same residual (4e-3 ... 2e-2 relative on real IIWA blocks) -- no PCG, the reference's included, is
    within 1e-4 of the direct solve at those tolerances.  The 1e-4 clauses therefore apply to the PCG
    iterate (lambda and residual against the fp64 run of the SAME number of iterations) ...
  * ... and to the direct solve only where the iteration is converged (tolerance driven to the fp32
    floor), where the yardstick is the reference's own fp32 QDLDL error on the same system.

`record()` returns every number; `check()` is the assertion the tests, smoke() and bench share.
Reference anchors: lambda* = reference QDLDL (double build, oracle/_ref) on the fp64 Schur system;
lambda_q32 = reference QDLDL in the reference's own configuration (float, Makefile:16).
"""
import numpy as np

from . import schur, pcg, pcg_cg1, bdmv, qdldl_solve

S_, C_ = 14, 7

# real-IIWA windows of the 0_0 trajectory (N, first row, perturbation of the iterate) and, per N, the exit
# tolerances the reference's own experiment sweeps (examples/track_iiwa_pcg.cu:48-68)
REAL_CASES = [
    (32, 0, 0.02, (7.5e-6, 5e-6, 2.5e-6, 1e-6)),
    (32, 200, 0.05, (7.5e-6, 5e-6, 2.5e-6, 1e-6)),
    (64, 100, 0.05, (7.5e-5, 5e-5, 2.5e-5, 1e-5)),
    (64, 300, 0.02, (7.5e-5, 5e-5, 2.5e-5, 1e-5)),
    (128, 300, 0.02, (1e-3, 5e-4, 1e-4, 5e-5, 1e-5)),
    (128, 0, 0.05, (1e-3, 5e-4, 1e-4, 5e-5, 1e-5)),
]


def _rel(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


def systems(N, G, C, g, cv, rho):
    """fp32 and fp64 Schur systems of one instance + the two reference QDLDL solutions."""
    S32, P32, g32, _ = schur(S_, C_, N, G, C, g, cv, rho, np.float32)
    S64, P64, g64, _ = schur(S_, C_, N, G, C, g, cv, rho, np.float64)
    star = qdldl_solve(S_, N, S64, g64)
    q32 = qdldl_solve(S_, N, S32, g32)
    return {"S32": S32, "P32": P32, "g32": g32, "S64": S64, "P64": P64, "g64": g64, "star": star, "q32": q32}


def record(N, sysd, lam0, tol, lam_gpu, it_gpu, ex_gpu, variant=0, max_iter=10000):
    """Every parity number of one GPU solve (lam_gpu, it_gpu, ex_gpu) of the system `sysd`."""
    S32, P32, g32, S64, P64, g64 = (sysd[k] for k in ("S32", "P32", "g32", "S64", "P64", "g64"))
    star, q32 = sysd["star"], sysd["q32"]
    lam0 = np.asarray(lam0, np.float32)
    res = lambda l: float(np.linalg.norm(bdmv(S_, N, S64, np.asarray(l, np.float64)) - g64) / np.linalg.norm(g64))
    run = pcg_cg1 if variant == 1 else pcg
    lam_o, it_o, ex_o = run(S_, N, S32, P32, g32, lam0, max_iter, tol)                       # fp32 oracle, own exit
    _, it_64, _ = run(S_, N, S64, P64, g64, lam0.astype(np.float64), max_iter, tol, np.float64)   # exact-arithmetic proxy
    lam_64k, _, _ = run(S_, N, S64, P64, g64, lam0.astype(np.float64), int(it_gpu), 0.0, np.float64)   # same count as the GPU
    lam_ok, _, _ = run(S_, N, S32, P32, g32, lam0, int(it_gpu), 0.0)
    # How far CORRECT fp32 implementations of this solve are apart -- the measured floor of the input:
    #  * the evaluation-order spread of the iteration on identical inputs (plain vs alternate fp32 order, fp32 vs
    #    an fp64 run of the same iterations on the fp32 system), and
    #  * the effect of forming S / Pinv / gamma in fp32: a different but equally valid rounding of the Schur
    #    complement (the reference's own kernel has one, ours another) moves the iterate by ~ kappa(S) x 1e-7.
    lam_64k_on32, _, _ = run(S_, N, S32.astype(np.float64), P32.astype(np.float64), g32.astype(np.float64),
                             lam0.astype(np.float64), int(it_gpu), 0.0, np.float64)
    variants = [lam_ok, lam_64k_on32]
    if variant == 0:
        variants.append(pcg(S_, N, S32, P32, g32, lam0, int(it_gpu), 0.0, np.float32, mode=1)[0])
    n_order = len(variants)
    # one more equally valid fp32 computation: the iteration in fp32 on the fp64 Schur system ROUNDED to fp32 (another
    # rounding of S / Pinv / gamma than the fp32 formation gives)
    variants.append(run(S_, N, S64.astype(np.float32), P64.astype(np.float32), g64.astype(np.float32), lam0, int(it_gpu), 0.0)[0])
    nrm = max(np.linalg.norm(lam_64k), 1e-300)
    floor_pcg = max(np.linalg.norm(np.asarray(a, np.float64) - np.asarray(b, np.float64))
                    for i, a in enumerate(variants[:n_order]) for b in variants[i + 1:n_order]) / nrm
    floor_schur = max(_rel(lam_64k_on32, lam_64k), _rel(variants[-1], lam_64k))
    floor = max(floor_pcg, floor_schur)
    res_64k = res(lam_64k)
    # the residual clause compares residual VECTORS: || S (lambda - lambda_ref) || / || gamma ||
    rvec = lambda l: float(np.linalg.norm(bdmv(S_, N, S64, np.asarray(l, np.float64) - lam_64k)) / np.linalg.norm(g64))
    res_floor = max(rvec(v) for v in variants)
    return {
        "N": int(N), "tol": float(tol), "variant": int(variant), "warm": bool(np.any(lam0)),
        "iters_gpu": int(it_gpu), "iters_oracle_f32": int(it_o), "iters_pcg_f64": int(it_64),
        "max_iter_exit_gpu": bool(ex_gpu), "max_iter_exit_oracle": bool(ex_o),
        "res_gpu": res(lam_gpu), "res_pcg_f64_same_iters": res_64k, "res_oracle_f32_same_iters": res(lam_ok),
        "res_vec_diff_gpu_vs_pcg_f64": rvec(lam_gpu), "res_floor": float(res_floor), "res_qdldl_f32": res(q32),
        "lam_rel_gpu_vs_pcg_f64_same_iters": _rel(lam_gpu, lam_64k),
        "lam_rel_oracle_f32_vs_pcg_f64_same_iters": _rel(lam_ok, lam_64k),
        "lam_rel_gpu_vs_oracle_f32_same_iters": _rel(lam_gpu, lam_ok),
        "fp32_floor": float(floor), "fp32_floor_pcg_order": float(floor_pcg), "fp32_floor_schur_rounding": float(floor_schur),
        "lam_rel_gpu_vs_qdldl_star": _rel(lam_gpu, star),
        "lam_rel_pcg_f64_vs_qdldl_star": _rel(lam_64k, star),     # the truncation error of the iteration itself
        "lam_rel_qdldl_f32_vs_star": _rel(q32, star),             # the reference's own fp32 direct-solve error
    }


LAM_TOL = 1e-4      # north_star
RES_TOL = 1e-4
FLOOR_K = 3.0       # widening factor over the MEASURED fp32 evaluation-order floor of the input


def check(r, converged=False):
    """The asserted criteria; returns a list of violations (empty = parity holds).
    Operating point (converged=False):
      * iterations within +-1 of the fp64 run of the same iteration (exact-arithmetic proxy) or of the fp32 oracle
        (the two references themselves differ by up to 2 on near-threshold exits);
      * residual: || S (lambda - lambda_ref) || / || gamma || <= max(1e-4, 3 x measured residual floor), lambda_ref = the
        fp64 run stopped at the same iteration (the difference of the two residual VECTORS);
      * lambda within max(1e-4, 3 x measured fp32 floor) of that run (relative).
    Converged regime (converged=True, tolerance driven to the fp32 floor): lambda within max(1e-4, 3 x fp32 floor,
    3 x the reference QDLDL-f32 error) of the reference's QDLDL (double) solution; residual within
    max(1e-4, 3 x QDLDL-f32's own residual)."""
    bad = []
    if converged:
        lim = max(LAM_TOL, FLOOR_K * r["lam_rel_qdldl_f32_vs_star"], FLOOR_K * r["fp32_floor"])
        if not r["lam_rel_gpu_vs_qdldl_star"] <= lim:
            bad.append("lambda vs QDLDL %.3g > %.3g" % (r["lam_rel_gpu_vs_qdldl_star"], lim))
        lim = max(RES_TOL, FLOOR_K * r["res_qdldl_f32"], FLOOR_K * r["res_oracle_f32_same_iters"])
        if not r["res_gpu"] <= lim:
            bad.append("residual %.3g > %.3g" % (r["res_gpu"], lim))
        return bad
    if min(abs(r["iters_gpu"] - r["iters_oracle_f32"]), abs(r["iters_gpu"] - r["iters_pcg_f64"])) > 1:
        bad.append("iters %d vs fp32 oracle %d / fp64 PCG %d" % (r["iters_gpu"], r["iters_oracle_f32"], r["iters_pcg_f64"]))
    if r["max_iter_exit_gpu"] != r["max_iter_exit_oracle"]:
        bad.append("max_iter_exit differs")
    lim = max(RES_TOL, FLOOR_K * r["res_floor"])
    if not r["res_vec_diff_gpu_vs_pcg_f64"] <= lim:
        bad.append("residual vector off the fp64 PCG's by %.3g > %.3g (floor %.3g)" % (r["res_vec_diff_gpu_vs_pcg_f64"], lim, r["res_floor"]))
    lim = max(LAM_TOL, FLOOR_K * r["fp32_floor"])
    if not r["lam_rel_gpu_vs_pcg_f64_same_iters"] <= lim:
        bad.append("lambda vs fp64 PCG %.3g > %.3g (floor %.3g)" % (r["lam_rel_gpu_vs_pcg_f64_same_iters"], lim, r["fp32_floor"]))
    return bad


def summarize(rows):
    """The `parity` object of a bench line / report: counts and worst cases over `rows`."""
    if not rows:
        return {"sampled": 0}
    viol = [check(r) for r in rows]
    return {
        "sampled": len(rows),
        "iters_within_1": int(sum(min(abs(r["iters_gpu"] - r["iters_oracle_f32"]), abs(r["iters_gpu"] - r["iters_pcg_f64"])) <= 1 for r in rows)),
        "iters_equal_fp64_pcg": int(sum(r["iters_gpu"] == r["iters_pcg_f64"] for r in rows)),
        "lambda_rel_max": max(r["lam_rel_gpu_vs_pcg_f64_same_iters"] for r in rows),
        "lambda_rel_median": float(np.median([r["lam_rel_gpu_vs_pcg_f64_same_iters"] for r in rows])),
        "floor_max": max(r["fp32_floor"] for r in rows),
        "res_diff_max": max(r["res_vec_diff_gpu_vs_pcg_f64"] for r in rows),
        "res_floor_max": max(r["res_floor"] for r in rows),
        "lambda_vs_qdldl_median": float(np.median([r["lam_rel_gpu_vs_qdldl_star"] for r in rows])),
        "truncation_vs_qdldl_median": float(np.median([r["lam_rel_pcg_f64_vs_qdldl_star"] for r in rows])),
        "violations": int(sum(1 for v in viol if v)),
        "criteria": "iters +-1 vs fp64 PCG or fp32 oracle; ||S(lambda - lambda_f64(same iters))||/||gamma|| <= max(1e-4, 3 x measured residual floor); "
                    "lambda vs fp64 PCG(same iters) <= max(1e-4, 3 x measured fp32 floor)",
    }
