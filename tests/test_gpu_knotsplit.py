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
