"""GPU parity tests (-m gpu): the sm_100a kernels, called through the C ABI, against the CPU
oracle on the same seeded inputs.  Tolerances (fp32, settings.cuh:49-57):
  * S / Pinv blocks, gamma, G^-1 : 1e-4 relative (Frobenius, per written block) vs the fp32
    oracle AND vs the fp64 oracle;
  * PCG lambda: 1e-4 relative vs the fp32 oracle PCG on identical (S, Pinv, gamma, lambda0);
    iteration count within +-1; residual within 1e-4 (relative to ||gamma||) of the oracle's;
  * vs the reference QDLDL direct solve: 1e-4 relative in the regime where fp32 CG can reach
    it (rho = 1: well conditioned); at rho = 1e-3 the comparison is reported, not asserted
    (kappa(S) makes two exact fp32 solvers disagree by > 1e-3).
"""
import os

import numpy as np
import pytest
import torch

import oracle
import mpcgpu_b200 as mg
from mpcgpu_b200 import synth
from helpers import S, C, rel, block_rel

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEV = "cuda:0"


def lam_tol(N, S32, P32, g32, lam0, iters):
    """1e-4 relative (north_star), widened to 3x the measured fp32 noise floor of THIS input:
    the distance between two correct fp32 evaluation orders / the fp64 run of the same
    iterations (oracle.pcg_noise_floor).  kappa(S) ~ 1e4..1e6 at rho = 1e-3 puts that floor
    between 2e-5 and 7e-4, see DESIGN.md "Parity"."""
    return max(1e-4, 3.0 * oracle.pcg_noise_floor(S, N, S32, P32, g32, lam0, iters))


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def gpu_schur(N, d, batch=1, rho=None):
    rho = d["rho"] if rho is None else rho
    z = mg.sizes(S, C, N)
    G, Cm, g, cv = dev(d["G"][:batch]), dev(d["C"][:batch]), dev(d["g"][:batch]), dev(d["c"][:batch])
    Sg = torch.full((batch, z["S"]), float("nan"), device=DEV)
    Pg = torch.full((batch, z["S"]), float("nan"), device=DEV)
    gam = torch.full((batch, z["gamma"]), float("nan"), device=DEV)
    mg.form_schur_system(S, C, N, G, Cm, g, cv, Sg, Pg, gam, rho, batch=batch)
    torch.cuda.synchronize()
    return Sg, Pg, gam, G


def gpu_pcg(N, Sg, Pg, gam, lam0, tol, max_iter=10000, batch=1, refstyle=False):
    lam = dev(lam0).clone()
    it = torch.full((batch,), -1, dtype=torch.int32, device=DEV)
    ex = torch.full((batch,), 7, dtype=torch.uint8, device=DEV)
    mg.pcg(S, N, Sg, Pg, gam, lam, it, ex, mg.PcgConfig(pcg_exit_tol=tol, pcg_max_iter=max_iter),
           batch=batch, refstyle=refstyle)
    torch.cuda.synchronize()
    return lam.cpu().numpy(), it.cpu().numpy(), ex.cpu().numpy()


@pytest.mark.parametrize("N", [1, 2, 3, 17, 32, 33, 64, 65, 128, 129, 256, 300, 512])
def test_schur_matches_oracle(N):
    d = synth.gen_kkt(N, 1, seed=synth.BASE_SEED + N)
    Sg, Pg, gam, Gi = gpu_schur(N, d)
    o32 = oracle.schur(S, C, N, d["G"][0], d["C"][0], d["g"][0], d["c"][0], d["rho"], np.float32)
    o64 = oracle.schur(S, C, N, d["G"][0], d["C"][0], d["g"][0], d["c"][0], d["rho"], np.float64)
    Sn, Pn, gn, Gn = (t.cpu().numpy()[0] for t in (Sg, Pg, gam, Gi))
    assert np.all(np.isfinite(Sn)) and np.all(np.isfinite(Pn))      # corners written as zeros
    for ref in (o32, o64):
        assert block_rel(S, N, Sn, ref[0]) < 1e-4
        assert block_rel(S, N, Pn, ref[1]) < 1e-4
        assert rel(gn, ref[2]) < 1e-4
        if N > 1:
            assert rel(Gn, ref[3]) < 1e-4
    z = Sn.reshape(N, 3, S * S)
    assert np.all(z[0, 0] == 0) and np.all(z[N - 1, 2] == 0)


@pytest.mark.parametrize("N,tol", [(1, 1e-8), (2, 1e-8), (16, 1e-6), (32, 1e-4), (33, 1e-6), (64, 1e-4),
                                   (64, 1e-6), (100, 1e-5), (128, 1e-4), (128, 1e-6), (256, 1e-4),
                                   (512, 1e-4)])
def test_pcg_matches_oracle(N, tol):
    d = synth.gen_kkt(N, 1, seed=synth.BASE_SEED + 3 * N)
    S32, P32, g32, _ = oracle.schur(S, C, N, d["G"][0], d["C"][0], d["g"][0], d["c"][0], d["rho"], np.float32)
    lam0 = np.zeros((1, S * N), np.float32)
    lam_o, it_o, ex_o = oracle.pcg(S, N, S32, P32, g32, lam0[0], 10000, tol)
    lam_g, it_g, ex_g = gpu_pcg(N, dev(S32[None]), dev(P32[None]), dev(g32[None]), lam0, tol)
    assert abs(int(it_g[0]) - it_o) <= 1, (it_g, it_o)
    assert bool(ex_g[0]) == ex_o
    # lambda / residual are compared at the SAME iteration count (a +-1 difference in the exit
    # iteration is allowed above and changes the residual by one CG step)
    it_cmp = int(it_g[0])
    lam_s, _, _ = oracle.pcg(S, N, S32, P32, g32, lam0[0], it_cmp, 0.0, np.float32, mode=0)
    assert rel(lam_g[0], lam_s) < lam_tol(N, S32, P32, g32, lam0[0], it_cmp)
    res = lambda l: np.linalg.norm(oracle.bdmv(S, N, S32.astype(np.float64), l.astype(np.float64)) - g32) / np.linalg.norm(g32)
    # residual within 1e-4 (relative to ||gamma||) of the oracle's, widened to 6x the spread
    # between the oracle's own two fp32 evaluation orders when the tolerance asks for more
    # than fp32 can deliver (tol 1e-8 runs into the fp32 stagnation floor)
    lam_alt, _, _ = oracle.pcg(S, N, S32, P32, g32, lam0[0], it_cmp, 0.0, np.float32, mode=1)
    # ... and, in that stagnation regime (the residual of ANY fp32 order is rounding noise of a few 1e-3 there and
    # differs by 2x between orders: N = 2 at tol 1e-8), "as converged as the oracle": within 3x of its two orders
    stagnated = res(lam_g[0]) < 3.0 * max(res(lam_s), res(lam_alt))
    assert abs(res(lam_g[0]) - res(lam_s)) < max(1e-4, 6.0 * abs(res(lam_alt) - res(lam_s))) or (tol <= 1e-8 and stagnated)


@pytest.mark.parametrize("N", [8, 64, 128])
def test_pcg_warm_start_and_max_iter(N):
    d = synth.gen_kkt(N, 1, seed=synth.BASE_SEED + 11 * N)
    S32, P32, g32, _ = oracle.schur(S, C, N, d["G"][0], d["C"][0], d["g"][0], d["c"][0], d["rho"], np.float32)
    S64, _, g64, _ = oracle.schur(S, C, N, d["G"][0], d["C"][0], d["g"][0], d["c"][0], d["rho"], np.float64)
    lam_star = oracle.qdldl_solve(S, N, S64, g64)
    rng = np.random.default_rng(N)
    warm = (lam_star * (1 + 1e-2 * rng.standard_normal(S * N))).astype(np.float32)[None]
    lam_o, it_o, ex_o = oracle.pcg(S, N, S32, P32, g32, warm[0], 10000, 1e-4)
    lam_g, it_g, ex_g = gpu_pcg(N, dev(S32[None]), dev(P32[None]), dev(g32[None]), warm, 1e-4)
    assert abs(int(it_g[0]) - it_o) <= 1 and bool(ex_g[0]) == ex_o
    assert rel(lam_g[0], lam_o) < lam_tol(N, S32, P32, g32, warm[0], it_o)
    # max_iter exhausted: iters == max_iter, flag = 1 (include/pcg/sqp.cuh:277,289)
    lam_o, it_o, ex_o = oracle.pcg(S, N, S32, P32, g32, np.zeros(S * N), 5, 1e-30)
    lam_g, it_g, ex_g = gpu_pcg(N, dev(S32[None]), dev(P32[None]), dev(g32[None]), np.zeros((1, S * N), np.float32), 1e-30, max_iter=5)
    assert it_o == 5 and int(it_g[0]) == 5 and ex_o and bool(ex_g[0])
    assert rel(lam_g[0], lam_o) < 1e-4


@pytest.mark.parametrize("N", [4, 64, 128])
def test_refstyle_kernel_matches_oracle(N):
    d = synth.gen_kkt(N, 1, seed=synth.BASE_SEED + 5 * N)
    S32, P32, g32, _ = oracle.schur(S, C, N, d["G"][0], d["C"][0], d["g"][0], d["c"][0], d["rho"], np.float32)
    lam_o, it_o, ex_o = oracle.pcg(S, N, S32, P32, g32, np.zeros(S * N), 10000, 1e-5)
    lam_g, it_g, ex_g = gpu_pcg(N, dev(S32[None]), dev(P32[None]), dev(g32[None]),
                                np.zeros((1, S * N), np.float32), 1e-5, refstyle=True)
    assert abs(int(it_g[0]) - it_o) <= 1 and bool(ex_g[0]) == ex_o
    assert rel(lam_g[0], lam_o) < lam_tol(N, S32, P32, g32, np.zeros(S * N, np.float32), it_o)


@pytest.mark.parametrize("N", [1, 5, 32, 64, 96, 128, 512])
def test_fused_matches_two_stage_and_oracle(N):
    B = 3
    d = synth.gen_kkt(N, B, seed=synth.BASE_SEED + 7 * N)
    tol = 1e-5
    cfg = mg.PcgConfig(pcg_exit_tol=tol)
    # two-stage on the GPU
    Sg, Pg, gam, Gi2 = gpu_schur(N, d, batch=B)
    lam2, it2, ex2 = gpu_pcg(N, Sg, Pg, gam, np.zeros((B, S * N), np.float32), tol, batch=B)
    # fused
    G, Cm, g, cv = dev(d["G"]), dev(d["C"]), dev(d["g"]), dev(d["c"])
    lam = torch.zeros((B, S * N), device=DEV)
    it = torch.zeros(B, dtype=torch.int32, device=DEV)
    ex = torch.zeros(B, dtype=torch.uint8, device=DEV)
    mg.kkt_solve(S, C, N, G, Cm, g, cv, d["rho"], lam, it, ex, cfg, batch=B)
    torch.cuda.synchronize()
    assert torch.equal(G, Gi2)                       # same G^-1, bit for bit
    assert np.array_equal(it.cpu().numpy(), it2) and np.array_equal(ex.cpu().numpy(), ex2)
    assert np.array_equal(lam.cpu().numpy(), lam2)   # same arithmetic, S/Pinv only kept on chip
    for b in range(B):
        S32, P32, g32, Gi = oracle.schur(S, C, N, d["G"][b], d["C"][b], d["g"][b], d["c"][b], d["rho"], np.float32)
        lam_o, it_o, ex_o = oracle.pcg(S, N, S32, P32, g32, np.zeros(S * N), 10000, tol)
        assert abs(int(it2[b]) - it_o) <= 1 and bool(ex2[b]) == ex_o
        # the GPU S differs from the oracle's S in the last bits (1e-6 relative); kappa(S)
        # amplifies that into lambda, exactly as it does between the oracle's own fp32 and
        # fp64 Schur systems.  Stage-wise parity (identical inputs) is asserted at 1e-4 above.
        S64, P64, g64, _ = oracle.schur(S, C, N, d["G"][b], d["C"][b], d["g"][b], d["c"][b], d["rho"], np.float64)
        lam_64, _, _ = oracle.pcg(S, N, S64, P64, g64, np.zeros(S * N), it_o, 0.0, np.float64)
        assert rel(lam2[b], lam_o) < max(1e-4, 4.0 * rel(lam_o, lam_64))


def test_pcg_reaches_reference_qdldl_when_well_conditioned():
    """north_star parity clause vs QDLDL: lambda and residual within 1e-4 relative."""
    N = 64
    d = synth.gen_kkt(N, 2)
    rho = 1.0
    Sg, Pg, gam, _ = gpu_schur(N, d, batch=2, rho=rho)
    lam, it, ex = gpu_pcg(N, Sg, Pg, gam, np.zeros((2, S * N), np.float32), 1e-12, batch=2)
    for b in range(2):
        Sn, gn = Sg[b].cpu().numpy(), gam[b].cpu().numpy()
        lq = oracle.qdldl_solve(S, N, Sn.astype(np.float64), gn.astype(np.float64))   # reference QDLDL
        lq32 = oracle.qdldl_solve(S, N, Sn, gn)                                        # its own fp32 config
        assert not ex[b]
        assert rel(lam[b], lq) < 1e-4
        assert rel(lam[b], lq32) < 1e-4
        assert rel(oracle.bdmv(S, N, Sn.astype(np.float64), lam[b].astype(np.float64)), gn) < 1e-4


def test_golden_fixture_through_gpu(golden_dir):
    z = np.load(os.path.join(golden_dir, "kkt_n8_warm.npz"))
    N = int(z["N"])
    d = {"G": z["G"][None], "C": z["C"][None], "g": z["g"][None], "c": z["cvec"][None], "rho": float(z["rho"])}
    Sg, Pg, gam, Gi = gpu_schur(N, d)
    assert block_rel(S, N, Sg[0].cpu().numpy(), z["S64"]) < 1e-4
    assert block_rel(S, N, Pg[0].cpu().numpy(), z["Pinv64"]) < 1e-4
    assert rel(Gi[0].cpu().numpy(), z["Ginv64"]) < 1e-4
    lam, it, ex = gpu_pcg(N, Sg, Pg, gam, z["lam0"][None], float(z["pcg_tol"]))
    assert abs(int(it[0]) - int(z["pcg_iters"])) <= 1
    # lambda against the committed oracle lambda: 1e-4, widened to 3 x the measured fp32 floor of this input (the
    # GPU's S differs from the fixture's in the last bits; kappa(S) amplifies that -- oracle/parity.py measures it)
    from oracle import parity
    sysd = parity.systems(N, z["G"], z["C"], z["g"], z["cvec"], float(z["rho"]))
    r = parity.record(N, sysd, z["lam0"], float(z["pcg_tol"]), lam[0], int(it[0]), bool(ex[0]))
    assert not parity.check(r), (parity.check(r), r["fp32_floor"])
    assert rel(lam[0], z["pcg_lambda_f32"]) < max(1e-4, 3.0 * r["fp32_floor"]), r["fp32_floor"]


def test_batched_full_size_properties():
    """BASELINE.json configs[4] scale (4096 x 64 knots): size-independent properties.
    (1) every instance's flags/iters are sane, (2) ||S lambda - gamma|| drops below the
    cold-start residual, (3) instance 0 equals the single-instance solve bit for bit,
    (4) permuting instances permutes the results (no cross-instance leakage)."""
    N, B = 64, 4096
    d = synth.gen_kkt_fast(N, B)
    cfg = mg.PcgConfig(pcg_exit_tol=1e-4)
    G, Cm, g, cv = dev(d["G"]), dev(d["C"]), dev(d["g"]), dev(d["c"])
    G0 = G.clone()
    lam = torch.zeros((B, S * N), device=DEV)
    it = torch.zeros(B, dtype=torch.int32, device=DEV)
    ex = torch.ones(B, dtype=torch.uint8, device=DEV)
    mg.kkt_solve(S, C, N, G, Cm, g, cv, d["rho"], lam, it, ex, cfg, batch=B)
    torch.cuda.synchronize()
    assert int(ex.sum()) == 0 and int(it.min()) >= 1 and int(it.max()) < 2000
    assert bool(torch.isfinite(lam).all())
    # residual through the materialised system
    z = mg.sizes(S, C, N)
    Sg = torch.empty((B, z["S"]), device=DEV); Pg = torch.empty_like(Sg); gam = torch.empty((B, S * N), device=DEV)
    G1 = G0.clone()
    mg.form_schur_system(S, C, N, G1, Cm, g, cv, Sg, Pg, gam, d["rho"], batch=B)
    torch.cuda.synchronize()
    assert torch.equal(G1, G)
    blk = Sg.view(B, N, 3, S, S).transpose(3, 4)          # [b][k][blk][row][col]
    l3 = lam.view(B, N, S)
    y = torch.einsum("bkrc,bkc->bkr", blk[:, :, 1], l3)
    y[:, 1:] += torch.einsum("bkrc,bkc->bkr", blk[:, 1:, 0], l3[:, :-1])
    y[:, :-1] += torch.einsum("bkrc,bkc->bkr", blk[:, :-1, 2], l3[:, 1:])
    res = (y - gam.view(B, N, S)).flatten(1).norm(dim=1) / gam.norm(dim=1)
    # every instance: the relative residual the oracle reaches on this workload at tol 1e-4 is 0.073 ... 0.088
    # (gpurun_out/parity_r2.jsonl, 32 sampled instances; the exit test is ABSOLUTE on r.Pinv r and ||gamma|| is
    # small here) -- all 4096 must land in that band; the per-instance criteria (iterations, residual and lambda
    # against the fp64 run, measured floors) are asserted on sampled instances in test_gpu_operating_point.py
    assert float(res.max()) < 0.2 and float(res.min()) > 0.03, (float(res.min()), float(res.max()))
    # single-instance equality + permutation
    perm = torch.randperm(B, device=DEV)
    G2 = G0[perm].contiguous()
    lam2 = torch.zeros_like(lam); it2 = torch.zeros_like(it); ex2 = torch.ones_like(ex)
    mg.kkt_solve(S, C, N, G2, Cm[perm].contiguous(), g[perm].contiguous(), cv[perm].contiguous(),
                 d["rho"], lam2, it2, ex2, cfg, batch=B)
    torch.cuda.synchronize()
    assert torch.equal(lam2, lam[perm]) and torch.equal(it2, it[perm])


def test_host_solver_matches_device_path():
    N, B = 64, 1500       # > one pipeline chunk
    d = synth.gen_kkt_fast(N, B, seed=99)
    cfg = mg.PcgConfig(pcg_exit_tol=1e-4)
    G, Cm, g, cv = dev(d["G"]), dev(d["C"]), dev(d["g"]), dev(d["c"])
    lam = torch.zeros((B, S * N), device=DEV)
    it = torch.zeros(B, dtype=torch.int32, device=DEV); ex = torch.ones(B, dtype=torch.uint8, device=DEV)
    mg.kkt_solve(S, C, N, G, Cm, g, cv, d["rho"], lam, it, ex, cfg, batch=B)
    torch.cuda.synchronize()
    hs = mg.HostSolver(S, C, N, B)
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    hG, hC, hg, hc = pin(d["G"]), pin(d["C"]), pin(d["g"]), pin(d["c"])
    hl = torch.zeros((B, S * N)).pin_memory()
    hi = torch.zeros(B, dtype=torch.int32).pin_memory(); he = torch.ones(B, dtype=torch.uint8).pin_memory()
    hGi = torch.zeros_like(hG).pin_memory()
    hs.solve(hG, hC, hg, hc, d["rho"], hl, hi, he, cfg, B, h_Ginv=hGi)
    assert torch.equal(hl, lam.cpu()) and torch.equal(hi, it.cpu()) and torch.equal(he, ex.cpu())
    assert torch.equal(hGi, G.cpu())
    # separate warm-start buffer, and the cold start (lambda0 = 0 without a copy): same answers
    hl2 = torch.full((B, S * N), float("nan")).pin_memory()
    hs.solve(hG, hC, hg, hc, d["rho"], hl2, hi, he, cfg, B, h_lambda0=torch.zeros((B, S * N)).pin_memory())
    assert torch.equal(hl2, lam.cpu())
    hl3 = torch.full((B, S * N), float("nan")).pin_memory()
    hs.solve(hG, hC, hg, hc, d["rho"], hl3, hi, he, cfg, B, h_lambda0=None)
    assert torch.equal(hl3, lam.cpu()) and torch.equal(hi, it.cpu())
    hs.close()


def test_errors_are_loud():
    with pytest.raises(mg.MpcGpuError):
        t = torch.zeros(10, device=DEV)
        mg.form_schur_system(12, 6, 4, t, t, t, t, t, t, t, 1e-3)       # unsupported block sizes
    with pytest.raises(mg.MpcGpuError):
        t = torch.zeros(10)
        mg.form_schur_system(S, C, 4, t, t, t, t, t, t, t, 1e-3)        # CPU tensors: no fallback


@pytest.mark.parametrize("N", [1, 2, 5, 32, 64, 96, 128, 512])
def test_compute_dz_matches_oracle_standalone_and_as_epilogue(N):
    """compute_dz (include/common/dz.cuh:3-133): standalone kernel against the oracle on identical
    (G^-1, C, g, lambda); the fused kernel's epilogue must give the standalone kernel's bits."""
    B = 3
    d = synth.gen_kkt(N, B, seed=synth.BASE_SEED + 11 * N)
    cfg = mg.PcgConfig(pcg_exit_tol=1e-5)
    G, Cm, g, cv = dev(d["G"]), dev(d["C"]), dev(d["g"]), dev(d["c"])
    lam = torch.zeros((B, S * N), device=DEV)
    it = torch.zeros(B, dtype=torch.int32, device=DEV)
    ex = torch.zeros(B, dtype=torch.uint8, device=DEV)
    nz = (S + C) * N - C
    dz_f = torch.full((B, nz), float("nan"), device=DEV)
    mg.kkt_solve_dz(S, C, N, G, Cm, g, cv, d["rho"], lam, dz_f, it, ex, cfg, batch=B)
    dz_s = torch.full((B, nz), float("nan"), device=DEV)
    mg.compute_dz(S, C, N, G, Cm, g, lam, dz_s, batch=B)   # G now holds G^-1
    torch.cuda.synchronize()
    assert torch.equal(dz_f, dz_s)
    Gi, lam_h, dz_h = G.cpu().numpy(), lam.cpu().numpy(), dz_s.cpu().numpy()
    assert np.all(np.isfinite(dz_h))
    for b in range(B):
        ref32 = oracle.dz(S, C, N, Gi[b], d["C"][b], d["g"][b], lam_h[b], np.float32)
        ref64 = oracle.dz(S, C, N, Gi[b], d["C"][b], d["g"][b], lam_h[b], np.float64)
        assert rel(dz_h[b], ref64) < max(1e-4, 4.0 * rel(ref32, ref64))


@pytest.mark.parametrize("N,B", [(1, 2), (5, 3), (33, 2), (64, 300), (100, 3), (512, 1)])
def test_kernels_write_only_their_outputs(N, B):
    """Every output buffer sits between two sentinel guards (and inputs are compared with their
    originals afterwards): nothing outside [buffer, buffer + size) may change, for ragged N (not a
    multiple of the knots per CTA), single-knot problems and batches larger than one wave."""
    z = mg.sizes(S, C, N)
    d = synth.gen_kkt_fast(N, B, seed=4242 + N)
    GUARD, SENT = 256, -12345.5

    def guarded(n, dtype=torch.float32, sent=SENT):
        buf = torch.full((n + 2 * GUARD,), sent, dtype=dtype, device=DEV)
        return buf, buf[GUARD:GUARD + n]

    def intact(buf, n, sent=SENT):
        return bool((buf[:GUARD] == sent).all()) and bool((buf[GUARD + n:] == sent).all())

    Gb, G = guarded(B * z["G"]); G.copy_(dev(d["G"]).reshape(-1))
    Cm, g, cv = dev(d["C"]), dev(d["g"]), dev(d["c"])
    C0, g0, c0 = Cm.clone(), g.clone(), cv.clone()
    Sb, Sg = guarded(B * z["S"]); Pb, Pg = guarded(B * z["S"]); gb, gam = guarded(B * z["gamma"])
    lb, lam = guarded(B * S * N); lam.zero_()
    db, dz = guarded(B * z["g"])
    ib, it = guarded(B, torch.int32, -7); eb, ex = guarded(B, torch.uint8, 77)
    cfg = mg.PcgConfig(pcg_exit_tol=1e-4)
    v = lambda t, n: t.view(B, n)
    mg.form_schur_system(S, C, N, v(G, z["G"]), Cm, g, cv, v(Sg, z["S"]), v(Pg, z["S"]), v(gam, z["gamma"]), d["rho"], batch=B)
    mg.pcg(S, N, v(Sg, z["S"]), v(Pg, z["S"]), v(gam, z["gamma"]), v(lam, S * N), it, ex, cfg, batch=B)
    mg.compute_dz(S, C, N, v(G, z["G"]), Cm, g, v(lam, S * N), v(dz, z["g"]), batch=B)
    G.copy_(dev(d["G"]).reshape(-1)); lam.zero_()
    mg.kkt_solve_dz(S, C, N, v(G, z["G"]), Cm, g, cv, d["rho"], v(lam, S * N), v(dz, z["g"]), it, ex, cfg, batch=B)
    torch.cuda.synchronize()
    assert intact(Gb, B * z["G"]) and intact(Sb, B * z["S"]) and intact(Pb, B * z["S"]) and intact(gb, B * z["gamma"])
    assert intact(lb, B * S * N) and intact(db, B * z["g"]) and intact(ib, B, -7) and intact(eb, B, 77)
    assert torch.equal(Cm, C0) and torch.equal(g, g0) and torch.equal(cv, c0)
    assert bool(torch.isfinite(lam).all()) and bool(torch.isfinite(dz).all()) and int(ex.sum()) == 0


@pytest.mark.parametrize("N,refstyle", [(8, False), (64, False), (128, False), (512, False), (64, True)])
def test_pcg_degenerate_exits_match_oracle(N, refstyle):
    """Exit paths of the iteration: max_iter = 0 (lambda untouched, 0 iterations, flag
    set), zero right-hand side with a zero start (p.Sp = 0 on the first iteration), a start that
    already satisfies |r.Pinv r| < tol."""
    d = synth.gen_kkt(N, 1, seed=synth.BASE_SEED + 13 * N)
    S32, P32, g32, _ = oracle.schur(S, C, N, d["G"][0], d["C"][0], d["g"][0], d["c"][0], d["rho"], np.float32)
    Sd, Pd, gd = dev(S32[None]), dev(P32[None]), dev(g32[None])
    rng = np.random.default_rng(7 * N)
    start = rng.standard_normal((1, S * N)).astype(np.float32)
    # max_iter = 0
    lam_o, it_o, ex_o = oracle.pcg(S, N, S32, P32, g32, start[0], 0, 1e-4)
    lam_g, it_g, ex_g = gpu_pcg(N, Sd, Pd, gd, start, 1e-4, max_iter=0, refstyle=refstyle)
    assert int(it_g[0]) == it_o == 0 and bool(ex_g[0]) == ex_o
    assert np.array_equal(lam_g[0], start[0]) and np.array_equal(lam_o, start[0])
    # gamma = 0, lambda0 = 0
    zero_g = dev(np.zeros((1, S * N), np.float32))
    lam_o, it_o, ex_o = oracle.pcg(S, N, S32, P32, 0 * g32, np.zeros(S * N, np.float32), 100, 1e-4)
    lam_g, it_g, ex_g = gpu_pcg(N, Sd, Pd, zero_g, np.zeros((1, S * N), np.float32), 1e-4, max_iter=100, refstyle=refstyle)
    assert int(it_g[0]) == it_o and bool(ex_g[0]) == ex_o
    assert np.array_equal(lam_g[0], np.zeros(S * N, np.float32)) and not np.any(lam_o)
    # converged start: one more solve from the converged lambda stops at once
    lam_c, _, _ = gpu_pcg(N, Sd, Pd, gd, np.zeros((1, S * N), np.float32), 1e-4, refstyle=refstyle)
    lam_o, it_o, ex_o = oracle.pcg(S, N, S32, P32, g32, lam_c[0], 100, 1e-2)
    lam_g, it_g, ex_g = gpu_pcg(N, Sd, Pd, gd, lam_c, 1e-2, max_iter=100, refstyle=refstyle)
    assert abs(int(it_g[0]) - it_o) <= 1 and it_o <= 1 and bool(ex_g[0]) == ex_o
    assert rel(lam_g[0], lam_o) < 1e-4
