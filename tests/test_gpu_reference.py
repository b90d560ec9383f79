"""GPU tests (-m gpu) against the REFERENCE'S OWN CUDA CODE.

oracle/_ref/run/ref_harness is compiled (oracle/build.py, in the build container) from the
reference's unchanged headers -- its form_schur_system kernel (include/pcg/linsys_setup.cuh), its
generate_kkt_submatrices with the GRiD IIWA dynamics (include/common/kkt.cuh) -- against the
drop-in shim headers of mpcgpu_b200/shim.  The prebuilt binary travels to the GPU box; nothing
here reads /root/reference at run time.

  * real IIWA KKT blocks (G, C, g, c) from the reference's KKT kernel on the 0_0 trajectory
  * S, Pinv, gamma, G^-1 from the reference's own Schur kernel  vs  our kernel  vs  the oracle
  * the shim's pcg<float,14,N> launched exactly like include/pcg/sqp.cuh:275  vs  the oracle
"""
import os
import subprocess

import numpy as np
import pytest
import torch

import oracle
import mpcgpu_b200 as mg
from helpers import S, C, rel, block_rel

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RUN = os.path.join(ROOT, "oracle", "_ref", "run")
HARNESS = os.path.join(RUN, "ref_harness")
DEV = "cuda:0"


def harness(*args):
    if not os.path.exists(HARNESS):
        pytest.fail("oracle/_ref/run/ref_harness missing: run oracle/build.py where /root/reference exists")
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = "/usr/local/cuda/lib64:" + env.get("LD_LIBRARY_PATH", "")
    r = subprocess.run([HARNESS, *[str(a) for a in args]], cwd=RUN, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, (r.returncode, r.stdout[-500:], r.stderr[-500:])


def split(buf, sizes):
    out, o = [], 0
    for n in sizes:
        out.append(buf[o:o + n]); o += n
    assert o == buf.size
    return out


def real_kkt(N, row0, perturb, tmp):
    z = mg.sizes(S, C, N)
    f = os.path.join(tmp, "kkt_%d_%d.bin" % (N, row0))
    harness("kkt", N, row0, perturb, f)
    G, Cm, g, cv = split(np.fromfile(f, np.float32), [z["G"], z["C"], z["g"], z["c"]])
    return f, G, Cm, g, cv


@pytest.mark.parametrize("N,row0,perturb", [(8, 0, 0.02), (32, 0, 0.02), (64, 100, 0.05), (128, 300, 0.02), (512, 0, 0.01)])
def test_schur_matches_reference_kernel_on_real_iiwa_blocks(N, row0, perturb, tmp_path):
    tmp = str(tmp_path)
    z = mg.sizes(S, C, N)
    rho = 1e-3
    fin, G, Cm, g, cv = real_kkt(N, row0, perturb, tmp)
    assert np.all(np.isfinite(G)) and np.all(np.isfinite(Cm)) and np.abs(cv[S:]).max() > 0
    fout = os.path.join(tmp, "schur.bin")
    harness("schur", N, rho, fin, fout)
    Sr, Pr, gr, Gir = split(np.fromfile(fout, np.float32), [z["S"], z["S"], z["gamma"], z["G"]])
    # ours
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(DEV)[None]
    Gd = t(G).clone()
    Sg = torch.full((1, z["S"]), float("nan"), device=DEV); Pg = torch.full_like(Sg, float("nan"))
    gam = torch.full((1, z["gamma"]), float("nan"), device=DEV)
    mg.form_schur_system(S, C, N, Gd, t(Cm), t(g), t(cv), Sg, Pg, gam, rho)
    torch.cuda.synchronize()
    So, Po, go, Gio = Sg[0].cpu().numpy(), Pg[0].cpu().numpy(), gam[0].cpu().numpy(), Gd[0].cpu().numpy()
    # oracle (fp32 and fp64 restatement)
    o32 = oracle.schur(S, C, N, G, Cm, g, cv, rho, np.float32)
    o64 = oracle.schur(S, C, N, G, Cm, g, cv, rho, np.float64)
    for name, (a, b) in {"ours-vs-reference": ((So, Po, go, Gio), (Sr, Pr, gr, Gir)),
                         "oracle32-vs-reference": (o32, (Sr, Pr, gr, Gir)),
                         "ours-vs-oracle64": ((So, Po, go, Gio), o64)}.items():
        assert block_rel(S, N, a[0], b[0]) < 1e-4, name
        assert block_rel(S, N, a[1], b[1]) < 1e-4, name
        assert rel(a[2], b[2]) < 1e-4, name
        assert rel(a[3], b[3]) < 1e-4, name
    # keep the captured blocks + reference outputs (small cases become committed golden fixtures)
    keep = os.path.join(ROOT, "gpurun_out")
    if N <= 32 and os.path.isdir(keep):
        np.savez_compressed(os.path.join(keep, "iiwa_kkt_n%d.npz" % N), N=N, row0=row0, perturb=perturb, rho=rho,
                            G=G, C=Cm, g=g, cvec=cv, S_ref=Sr, Pinv_ref=Pr, gamma_ref=gr, Ginv_ref=Gir)


@pytest.mark.parametrize("N,row0", [(8, 0), (32, 50), (64, 100), (128, 300), (512, 0)])
def test_l1_shim_stage1_is_our_schur_kernel(N, row0, tmp_path):
    """Stage 1 behind the L1 shim (VERDICT r1 "Next" #4): `ref_harness schur_shim` calls the function the unchanged
    include/pcg/sqp.cuh:243 binds to under the shim (the form_schur_system<float> specialisation of
    mpcgpu_b200/shim/gpu_pcg.cuh -> mpcb200::schur_kernel).  Its S / Pinv / gamma / G^-1 must be the bits of the C ABI's
    two-stage call (same kernel, same tiling) and within 1e-4 of the reference's own kernel on the same real IIWA blocks."""
    tmp = str(tmp_path)
    z = mg.sizes(S, C, N)
    rho = 1e-3
    fin, G, Cm, g, cv = real_kkt(N, row0, 0.03, tmp)
    f_ref, f_shim = os.path.join(tmp, "schur_ref.bin"), os.path.join(tmp, "schur_shim.bin")
    harness("schur", N, rho, fin, f_ref)
    harness("schur_shim", N, rho, fin, f_shim)
    ref = split(np.fromfile(f_ref, np.float32), [z["S"], z["S"], z["gamma"], z["G"]])
    shim = split(np.fromfile(f_shim, np.float32), [z["S"], z["S"], z["gamma"], z["G"]])
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(DEV)[None]
    Gd = t(G).clone()
    Sg = torch.zeros((1, z["S"]), device=DEV); Pg = torch.zeros_like(Sg); gam = torch.zeros((1, z["gamma"]), device=DEV)
    mg.form_schur_system(S, C, N, Gd, t(Cm), t(g), t(cv), Sg, Pg, gam, rho)
    torch.cuda.synchronize()
    ours = (Sg[0].cpu().numpy(), Pg[0].cpu().numpy(), gam[0].cpu().numpy(), Gd[0].cpu().numpy())
    blk = lambda a: a.reshape(N, 3, S * S)
    for a, b, is_bd in zip(shim, ours, (True, True, False, False)):
        if is_bd:      # the never-written corner blocks hold whatever the caller put there (the harness zero-fills, as we do)
            assert np.array_equal(blk(a), blk(b))
        else:
            assert np.array_equal(a, b)
    assert block_rel(S, N, shim[0], ref[0]) < 1e-4 and block_rel(S, N, shim[1], ref[1]) < 1e-4
    assert rel(shim[2], ref[2]) < 1e-4 and rel(shim[3], ref[3]) < 1e-4


@pytest.mark.parametrize("N,tol", [(8, 1e-6), (32, 1e-5), (64, 1e-4), (128, 1e-4), (512, 1e-4)])
def test_shim_pcg_launched_like_the_reference_slot(N, tol, tmp_path):
    """pcg<float,14,N> of mpcgpu_b200/shim/gpu_pcg.cuh behind the reference's launch
    (cooperative, grid = N, 128 threads, pcgSharedMemSize; include/pcg/sqp.cuh:275-277)."""
    tmp = str(tmp_path)
    z = mg.sizes(S, C, N)
    _, G, Cm, g, cv = real_kkt(N, 50, 0.03, tmp)
    S32, P32, g32, _ = oracle.schur(S, C, N, G, Cm, g, cv, 1e-3, np.float32)
    lam0 = np.zeros(S * N, np.float32)
    fin, fout = os.path.join(tmp, "pcg_in.bin"), os.path.join(tmp, "pcg_out.bin")
    np.concatenate([S32, P32, g32, lam0]).astype(np.float32).tofile(fin)
    harness("pcg", N, tol, 10000, fin, fout)
    o = np.fromfile(fout, np.float32)
    lam_g, it_g, ex_g = o[:S * N], int(o[S * N]), bool(o[S * N + 1])
    lam_o, it_o, ex_o = oracle.pcg(S, N, S32, P32, g32, lam0, 10000, tol)
    assert abs(it_g - it_o) <= 1 and ex_g == ex_o, (it_g, it_o)
    lam_s, _, _ = oracle.pcg(S, N, S32, P32, g32, lam0, it_g, 0.0)
    assert rel(lam_g, lam_s) < max(1e-4, 3.0 * oracle.pcg_noise_floor(S, N, S32, P32, g32, lam0, it_g))


def _track(exe_name, tol, env_extra=None):
    import json
    exe = os.path.join(RUN, exe_name)
    if not os.path.exists(exe):
        pytest.fail("oracle/_ref/run/%s missing: run oracle/build.py where /root/reference exists" % exe_name)
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = "/usr/local/cuda/lib64:" + env.get("LD_LIBRARY_PATH", "")
    env.update(env_extra or {})
    r = subprocess.run([exe, str(tol)], cwd=RUN, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (exe_name, r.stdout[-400:], r.stderr[-400:])
    out = json.loads(r.stdout.strip().splitlines()[-1])
    keep = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(keep):
        with open(os.path.join(keep, "closed_loop_tests.jsonl"), "a") as f:
            f.write(json.dumps(dict(out, exe=exe_name, env=env_extra or {})) + "\n")
    return out


@pytest.mark.parametrize("N,tol", [(32, 5e-6), (128, 1e-4)])
def test_closed_loop_tracking_through_the_slots(N, tol):
    """The reference's unchanged simulateMPC (include/mpcsim.cuh, 2 ms SQP budget per control step, 0_0 figure
    eight; tools/track_slot.cu) over four bindings of its solver slot, same box, same run:
      ref   include/pcg/sqp.cuh with the reference's OWN form_schur_system kernel + the shim's pcg (round-1 L1)
      l1    include/pcg/sqp.cuh on the whole L1 shim (stage 1 bound to mpcb200::schur_kernel as well)
      b200  mpcgpu_b200/shim/b200/sqp.cuh (fused solve, graphed iteration), uncapped
      cap   b200 limited to the SQP iterations per step the reference slot reaches (MPCGPU_B200_SLOT_MAX_SQP)
    Measured (profiles/r2a_track_capped.jsonl): at N = 32 the tracking error of THIS MPC formulation grows with the
    SQP iterations spent per control step (0.075 / 0.104 / 0.114 / 0.118 at 2 / 4 / 8 / 16 iterations) whichever
    slot runs them, so the slots are compared at EQUAL iteration budget, and the uncapped run only against the
    measured trend."""
    import math
    ref = _track("track_slot_pcgrefschur_n%d" % N, tol)
    l1 = _track("track_slot_pcg_n%d" % N, tol)
    b = _track("track_slot_b200_n%d" % N, tol)
    cap = max(1, int(math.ceil(ref["sqp_iters_mean"])))
    bc = _track("track_slot_b200_n%d" % N, tol, {"MPCGPU_B200_SLOT_MAX_SQP": str(cap)})
    for o in (ref, l1, b, bc):
        assert o["control_steps"] > 1000 and np.isfinite(o["tracking_err_mean"])
    # same iteration budget -> the b200 slot tracks at least as well as the reference slot (it completes every
    # iteration it starts; the reference slot's budget often expires between its stages)
    assert bc["sqp_iters_mean"] <= cap + 1e-6
    assert bc["tracking_err_mean"] <= 1.25 * ref["tracking_err_mean"], (bc, ref)
    # the slot's purpose: several times the reference slot's SQP iterations inside the same 2 ms
    assert b["sqp_iters_mean"] >= 2.0 * ref["sqp_iters_mean"], (b, ref)
    assert l1["sqp_iters_mean"] >= 0.9 * ref["sqp_iters_mean"], (l1, ref)
    assert l1["tracking_err_mean"] <= 1.35 * ref["tracking_err_mean"], (l1, ref)
    if N >= 128:
        assert b["tracking_err_mean"] <= 1.10 * ref["tracking_err_mean"], (b, ref)     # measured 0.058-0.064 vs 0.078
    else:
        assert b["tracking_err_mean"] <= 1.35 * ref["tracking_err_mean"], (b, ref)     # measured 0.118 vs 0.098 (see docstring)


@pytest.mark.parametrize("N,tol", [(32, 5e-6), (128, 1e-4)])
def test_one_sqp_iteration_of_either_slot_on_identical_state(N, tol):
    """tools/slot_ab.cu: ONE SQP iteration of the reference's unchanged PCG slot (its own Schur kernel, its compute_dz,
    its line search, cuBLAS update) against ONE iteration of the b200 slot from identical (xu, lambda0, rho), 6 chained
    iterations on 4 windows of the 0_0 trajectory (warm lambda, adapting rho): same line-search decision and rho every
    time, PCG iteration counts within +-1 (one near-threshold exception allowed per 24), the applied step alpha dz
    within the fp32 floor of the window (median <= 1e-3)."""
    import json
    exe = os.path.join(RUN, "slot_ab_n%d" % N)
    if not os.path.exists(exe):
        pytest.fail("oracle/_ref/run/slot_ab_n%d missing: run oracle/build.py where /root/reference exists" % N)
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = "/usr/local/cuda/lib64:" + env.get("LD_LIBRARY_PATH", "")
    r = subprocess.run([exe, str(tol), "6"], cwd=RUN, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout[-400:], r.stderr[-400:])
    out = json.loads(r.stdout.strip().splitlines()[-1])
    n = out["sqp_iterations_compared"]
    assert n == 24
    assert out["same_rho"] == n and out["same_accept_decision"] == n
    assert out["pcg_iters_within_1"] >= n - 1 and out["max_pcg_iter_diff"] <= 4
    steps = [p["step_rel"] for p in out["per_iteration"] if p["accepted"] == [1, 1]]
    assert out["same_step_length"] >= out["accepted_steps"] - 1          # a near-tie of two merits may pick the other length
    assert float(np.median(steps)) <= 1e-3 and float(np.median([p["lambda_rel"] for p in out["per_iteration"]])) <= 1e-3


@pytest.mark.parametrize("N,B", [(8, 5), (32, 40), (64, 300), (128, 64)])
def test_batched_kkt_producer_matches_reference_kernel_and_feeds_batched_solve(N, B):
    """SURVEY 8(f) rank 3 / BASELINE.json configs[4] end to end on real IIWA dynamics
    (tools/kkt_batched_check.cu): b200slot::generate_kkt_submatrices_batched writes, for every one of
    B perturbed instances, the bits the reference's own generate_kkt_submatrices
    (include/common/kkt.cuh:22-163) writes for that instance alone; the batched kkt_solve_dz on those
    blocks equals single-instance solves bit for bit and converges on every instance."""
    import json
    exe = os.path.join(RUN, "kkt_batched_check")
    if not os.path.exists(exe):
        pytest.fail("oracle/_ref/run/kkt_batched_check missing: run oracle/build.py where /root/reference exists")
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = "/usr/local/cuda/lib64:" + env.get("LD_LIBRARY_PATH", "")
    r = subprocess.run([exe, str(N), str(B), "1e-4"], cwd=RUN, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, (r.returncode, r.stdout[-800:], r.stderr[-500:])
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["kkt_mismatches_vs_reference_kernel"] == 0 and out["solve_mismatches_vs_single"] == 0
    assert out["nonfinite_dz"] == 0 and out["max_iter_exits"] == 0 and out["pcg_iters"]["min"] >= 1


@pytest.mark.parametrize("N,B,K", [(8, 6, 3), (32, 128, 4), (64, 200, 3)])
def test_batched_sqp_iterations_follow_the_reference_sequence(N, B, K):
    """SURVEY 8(f) rank 4 (tools/sqp_batched_check.cu): K SQP iterations of B instances with no host in
    the loop (batched linearisation, fused solve with per-instance rho, all 8 line-search step lengths of
    all instances in one launch, device-side step selection / rho adaptation / update) end, for every
    checked instance, on the bits the reference's own per-instance sequence produces
    (generate_kkt_submatrices, 8 cooperative ls_gato_compute_merit launches, the host decision of
    include/pcg/sqp.cuh:352-413)."""
    import json
    exe = os.path.join(RUN, "sqp_batched_check")
    if not os.path.exists(exe):
        pytest.fail("oracle/_ref/run/sqp_batched_check missing: run oracle/build.py where /root/reference exists")
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = "/usr/local/cuda/lib64:" + env.get("LD_LIBRARY_PATH", "")
    r = subprocess.run([exe, str(N), str(B), str(K), "1e-4"], cwd=RUN, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, (r.returncode, r.stdout[-900:], r.stderr[-500:])
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["xu_mismatches_vs_reference_sequence"] == 0 and out["rho_mismatches"] == 0 and out["merit_mismatches"] == 0
    assert out["nonfinite_xu"] == 0
    if N >= 32:     # (the 8-knot windows start at their optimum: no step length lowers the merit, rho is raised)
        assert out["instances_with_lower_merit"] == B
