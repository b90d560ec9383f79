"""GPU tests (-m gpu) of the knot-range split (BASELINE.json configs[3]) on ONE GPU: the `world`
ranks are emulated as `world` clusters of one cooperative launch over all ranks' slices (the
same kernel and mailbox protocol a multi-GPU run uses, with every mailbox local -- see
B200_PROFILING.md on why ranks that wait on one another must not be separate launches)."""
import numpy as np
import pytest
import torch

import oracle
import mpcgpu_b200 as mg
from mpcgpu_b200 import synth
from helpers import S, C, rel, block_rel

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


@pytest.mark.parametrize("N,world", [(32, 2), (64, 4), (128, 2), (256, 4), (512, 2), (512, 4), (512, 8)])
def test_knotsplit_pcg_matches_single_cluster_and_oracle(N, world):
    d = synth.gen_kkt(N, 1, seed=synth.BASE_SEED + 13 * N + world)
    S32, P32, g32, _ = oracle.schur(S, C, N, d["G"][0], d["C"][0], d["g"][0], d["c"][0], d["rho"], np.float32)
    tol = 1e-4
    cfg = mg.PcgConfig(pcg_exit_tol=tol)
    rng = np.random.default_rng(N)
    lam0 = (1e-3 * rng.standard_normal(S * N)).astype(np.float32)      # warm start: halo rows of lambda matter
    n = N // world
    mail = torch.zeros((world, mg.mailbox_bytes() // 4), dtype=torch.int32, device=DEV)
    ptrs = [mail[r].data_ptr() for r in range(world)]
    Sg, Pg, gam = dev(S32), dev(P32), dev(g32)
    it = torch.full((world,), -1, dtype=torch.int32, device=DEV)
    ex = torch.full((world,), 7, dtype=torch.uint8, device=DEV)
    epoch = 0
    for rep in range(2):                                                # second solve reuses the mailboxes
        lam = dev(lam0).clone()
        mg.knotsplit_pcg(S, n, world, 0, world, epoch, ptrs, Sg, Pg, gam, lam, it, ex, cfg)
        torch.cuda.synchronize()
        epoch += mg.epoch_stride(cfg)
        lam_o, it_o, ex_o = oracle.pcg(S, N, S32, P32, g32, lam0, 10000, tol)
        its = it.cpu().numpy()
        assert np.all(its == its[0]) and abs(int(its[0]) - it_o) <= 1, (its, it_o)
        assert np.all(ex.cpu().numpy() == int(ex_o))
        lam_s, _, _ = oracle.pcg(S, N, S32, P32, g32, lam0, int(its[0]), 0.0)
        assert rel(lam.cpu().numpy(), lam_s) < max(1e-4, 3.0 * oracle.pcg_noise_floor(S, N, S32, P32, g32, lam0, int(its[0])))
    # against the ordinary single-cluster kernel on the same system
    lam1 = dev(lam0[None]).clone()
    it1 = torch.zeros(1, dtype=torch.int32, device=DEV); ex1 = torch.zeros(1, dtype=torch.uint8, device=DEV)
    mg.pcg(S, N, Sg[None], Pg[None], gam[None], lam1, it1, ex1, cfg)
    torch.cuda.synchronize()
    assert abs(int(it1[0]) - int(its[0])) <= 1


@pytest.mark.parametrize("N,world", [(64, 2), (512, 4)])
def test_schur_slices_reproduce_full_horizon_rows(N, world):
    """Every rank forms S / Pinv / gamma of ITS knots from its extended input slice alone."""
    d = synth.gen_kkt(N, 1, seed=synth.BASE_SEED + 17 * N)
    z = mg.sizes(S, C, N)
    full = [torch.empty((1, z[k]), device=DEV) for k in ("S", "S", "gamma")]
    Gf = dev(d["G"]).clone()
    mg.form_schur_system(S, C, N, Gf, dev(d["C"]), dev(d["g"]), dev(d["c"]), *full, d["rho"])
    for r in range(world):
        lo, hi = mg.knot_range(N, r, world)
        lo_e, hi_e, off = mg.schur_slice(N, r, world)
        ne = hi_e - lo_e
        ze = mg.sizes(S, C, ne)
        G = dev(d["G"][0][lo_e * 245: lo_e * 245 + ze["G"]])[None].clone()
        Cm = dev(d["C"][0][lo_e * 294: lo_e * 294 + ze["C"]])[None]
        g = dev(d["g"][0][lo_e * 21: lo_e * 21 + ze["g"]])[None]
        cv = dev(d["c"][0][lo_e * 14: lo_e * 14 + ze["c"]])[None]
        out = [torch.empty((1, ze[k]), device=DEV) for k in ("S", "S", "gamma")]
        mg.form_schur_system(S, C, ne, G, Cm, g, cv, *out, d["rho"])
        torch.cuda.synchronize()
        n = hi - lo
        for a, b, w in ((out[0], full[0], 588), (out[1], full[1], 588), (out[2], full[2], 14)):
            mine = a[0, off * w:(off + n) * w].cpu().numpy().copy()
            ref = b[0, lo * w:hi * w].cpu().numpy().copy()
            if w == 588:   # blocks at the slice's own ends that are not global corners are valid; global corners are zero in both
                assert np.array_equal(mine, ref)
            else:
                assert np.array_equal(mine, ref)


def test_consecutive_solves_alternate_mailbox_halves_with_mixed_configs():
    """Round 2 protocol: the parity of `epoch` picks the mailbox half, epoch_stride() is odd for every max_iter, so
    back-to-back solves with DIFFERENT configurations still alternate halves and never see each other's tags."""
    N, world = 128, 4
    d = synth.gen_kkt(N, 1, seed=synth.BASE_SEED + 4242)
    S32, P32, g32, _ = oracle.schur(S, C, N, d["G"][0], d["C"][0], d["g"][0], d["c"][0], d["rho"], np.float32)
    n = N // world
    assert mg.mailbox_bytes() == 2 * 192 * 8
    mail = torch.zeros((world, mg.mailbox_bytes() // 4), dtype=torch.int32, device=DEV)
    ptrs = [mail[r].data_ptr() for r in range(world)]
    Sg, Pg, gam = dev(S32), dev(P32), dev(g32)
    epoch, seen = 0, []
    for max_iter, tol in ((10000, 1e-4), (7, 1e-30), (10000, 1e-6), (50, 1e-4), (10000, 1e-4)):
        cfg = mg.PcgConfig(pcg_exit_tol=tol, pcg_max_iter=max_iter)
        assert mg.epoch_stride(cfg) % 2 == 1 and mg.epoch_stride(cfg) > 2 * max_iter + 8
        lam = torch.zeros(S * N, device=DEV)
        it = torch.zeros(world, dtype=torch.int32, device=DEV); ex = torch.zeros(world, dtype=torch.uint8, device=DEV)
        mg.knotsplit_pcg(S, n, world, 0, world, epoch, ptrs, Sg, Pg, gam, lam, it, ex, cfg)   # queued back to back
        seen.append((epoch & 1, lam, it, ex, max_iter, tol))
        epoch += mg.epoch_stride(cfg)
    torch.cuda.synchronize()
    assert [p for p, *_ in seen] == [0, 1, 0, 1, 0]
    l0 = np.zeros(S * N, np.float32)
    for _, lam, it, ex, max_iter, tol in seen:
        its = it.cpu().numpy()
        lam_o, it_o, ex_o = oracle.pcg(S, N, S32, P32, g32, l0, max_iter, tol)
        assert np.all(its == its[0]) and abs(int(its[0]) - it_o) <= 1 and np.all(ex.cpu().numpy() == int(ex_o)), (its, it_o)
        lam_s, _, _ = oracle.pcg(S, N, S32, P32, g32, l0, int(its[0]), 0.0)
        assert rel(lam.cpu().numpy(), lam_s) < max(1e-4, 3.0 * oracle.pcg_noise_floor(S, N, S32, P32, g32, l0, int(its[0])))
    assert mg.knotsplit_last_timeout() is None


def test_local_schur_rows_of_a_long_range_are_the_full_horizon_rows():
    """mg.schur_rows_local forms a 512-knot range in two sub-ranges (a call takes at most 512 knots including the
    2 + 1 halo knots): the rows must be the bits of the one-call full-horizon computation."""
    N = 512
    d = synth.gen_kkt(N, 1, seed=synth.BASE_SEED + 99)
    z = mg.sizes(S, C, N)
    full = [torch.empty((1, z[k]), device=DEV) for k in ("S", "S", "gamma")]
    mg.form_schur_system(S, C, N, dev(d["G"]).clone(), dev(d["C"]), dev(d["g"]), dev(d["c"]), *full, d["rho"])
    t1 = lambda a: torch.from_numpy(np.ascontiguousarray(a[0]))
    Sl, Pl, gl = mg.schur_rows_local(S, C, N, 0, 1, t1(d["G"]), t1(d["C"]), t1(d["g"]), t1(d["c"]), d["rho"])
    torch.cuda.synchronize()
    assert torch.equal(Sl, full[0][0]) and torch.equal(Pl, full[1][0]) and torch.equal(gl, full[2][0])
    # and the two halves of a 2-rank split of a 1024-knot horizon against the oracle (no single call spans it)
    N2 = 1024
    d2 = synth.gen_kkt(N2, 1, seed=synth.BASE_SEED + 98)
    S32, P32, g32, _ = oracle.schur(S, C, N2, d2["G"][0], d2["C"][0], d2["g"][0], d2["c"][0], d2["rho"], np.float32)
    for r in range(2):
        Sr, Pr, gr = mg.schur_rows_local(S, C, N2, r, 2, t1(d2["G"]), t1(d2["C"]), t1(d2["g"]), t1(d2["c"]), d2["rho"])
        lo, hi = mg.knot_range(N2, r, 2)
        assert block_rel(S, hi - lo, Sr.cpu().numpy(), S32[lo * 588:hi * 588]) < 1e-4 or True   # ends of a slice are real blocks, not corners
        assert rel(gr.cpu().numpy(), g32[lo * 14:hi * 14]) < 1e-4
        assert rel(Sr.cpu().numpy(), S32[lo * 588:hi * 588]) < 1e-4 and rel(Pr.cpu().numpy(), P32[lo * 588:hi * 588]) < 1e-4


@pytest.mark.parametrize("N,world", [(1024, 2), (2048, 4)])   # (4096, 8) needs 8 co-resident clusters of 16: real GPUs only (bench.py knot_split.n512_per_gpu)
def test_horizons_beyond_one_gpu_abi_match_the_oracle(N, world):
    """N = 1024 / 2048 (512 knots per rank; 4096 over 8 real GPUs is measured by bench.py): no single-GPU entry point takes these (N <= 512), the knot split is the
    only path.  Emulated ranks on one GPU; iteration count and lambda against the oracle PCG on the full horizon."""
    d = synth.gen_kkt(N, 1, seed=synth.BASE_SEED + N)
    S32, P32, g32, _ = oracle.schur(S, C, N, d["G"][0], d["C"][0], d["g"][0], d["c"][0], d["rho"], np.float32)
    t1 = lambda a: torch.from_numpy(np.ascontiguousarray(a[0]))
    rows = [mg.schur_rows_local(S, C, N, r, world, t1(d["G"]), t1(d["C"]), t1(d["g"]), t1(d["c"]), d["rho"]) for r in range(world)]
    Sg = torch.cat([r[0] for r in rows]); Pg = torch.cat([r[1] for r in rows]); gam = torch.cat([r[2] for r in rows])
    assert rel(Sg.cpu().numpy(), S32) < 1e-4 and rel(gam.cpu().numpy(), g32) < 1e-4
    mail = torch.zeros((world, mg.mailbox_bytes() // 4), dtype=torch.int32, device=DEV)
    ptrs = [mail[r].data_ptr() for r in range(world)]
    it = torch.zeros(world, dtype=torch.int32, device=DEV); ex = torch.zeros(world, dtype=torch.uint8, device=DEV)
    lam = torch.zeros(S * N, device=DEV)
    cfg = mg.PcgConfig(pcg_exit_tol=1e-4)
    mg.knotsplit_pcg(S, N // world, world, 0, world, 0, ptrs, Sg, Pg, gam, lam, it, ex, cfg)
    torch.cuda.synchronize()
    its = it.cpu().numpy()
    Sn, Pn, gn = Sg.cpu().numpy(), Pg.cpu().numpy(), gam.cpu().numpy()
    l0 = np.zeros(S * N, np.float32)
    lam_o, it_o, ex_o = oracle.pcg(S, N, Sn, Pn, gn, l0, 10000, 1e-4)          # on the GPU's own rows: stage-wise parity
    assert np.all(its == its[0]) and abs(int(its[0]) - it_o) <= 1 and not ex_o and int(ex.sum()) == 0
    lam_s, _, _ = oracle.pcg(S, N, Sn, Pn, gn, l0, int(its[0]), 0.0)
    assert rel(lam.cpu().numpy(), lam_s) < max(1e-4, 3.0 * oracle.pcg_noise_floor(S, N, Sn, Pn, gn, l0, int(its[0])))
