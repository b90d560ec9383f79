"""GPU stress tests (-m gpu) standing in for compute-sanitizer racecheck / synccheck, which is CLOSED on this pool
(profiles/r2_compute_sanitizer_closed_on_pool.txt: exit 86).  A missing barrier or an unordered st.async / DSMEM / mailbox
access shows up as run-to-run differences: every protocol of the library (CL = 1 with tensor memory, clusters of 2 / 4 / 8 /
16 with st.async + mbarrier reductions and halo pushes, the Pr scatter into the left neighbour's rows, the fp64 twins, the
knot-range split's tagged mailboxes) is run many times under changing co-scheduling (other launches of different shapes
in flight on other streams, batches that do and do not fill the machine) and must give the SAME BITS every time, equal
to the single-instance result.  Out-of-bounds writes are covered by test_kernels_write_only_their_outputs (guard bands)."""
import numpy as np
import pytest
import torch

import oracle
import mpcgpu_b200 as mg
from mpcgpu_b200 import synth
from helpers import S, C

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def dev(a, dt=None):
    x = torch.from_numpy(np.ascontiguousarray(a))
    return (x.to(dt) if dt else x).to(DEV)


def _noise(stream, d, N, B):
    """Another batch of another shape on another stream: perturbs which SMs / clusters are free."""
    with torch.cuda.stream(stream):
        G = dev(d["G"][:B]).clone(); lam = torch.zeros((B, S * N), device=DEV)
        it = torch.zeros(B, dtype=torch.int32, device=DEV); ex = torch.zeros(B, dtype=torch.uint8, device=DEV)
        mg.kkt_solve(S, C, N, G, dev(d["C"][:B]), dev(d["g"][:B]), dev(d["c"][:B]), d["rho"], lam, it, ex,
                     mg.PcgConfig(pcg_exit_tol=1e-4, pcg_max_iter=40), batch=B)
    return lam


@pytest.mark.parametrize("N,B,latency,dtype", [(64, 300, 0, torch.float32), (64, 2, 1, torch.float32), (100, 40, 0, torch.float32),
                                               (128, 1, 1, torch.float32), (256, 1, 1, torch.float32), (512, 1, 1, torch.float32),
                                               (512, 9, 0, torch.float32), (64, 2, 1, torch.float64), (128, 30, 0, torch.float64)])
def test_repeated_solves_give_the_same_bits_under_changing_co_scheduling(N, B, latency, dtype, monkeypatch):
    monkeypatch.setenv("MPCGPU_B200_LATENCY_TILING", str(latency))
    d = synth.gen_kkt_fast(N, B, seed=500 + N)
    dn = synth.gen_kkt_fast(32, 64, seed=9)
    G0, Cm, g, cv = dev(d["G"], dtype), dev(d["C"], dtype), dev(d["g"], dtype), dev(d["c"], dtype)
    nz = (S + C) * N - C
    cfg = mg.PcgConfig(pcg_exit_tol=1e-5)
    side = torch.cuda.Stream()
    ref = None
    for rep in range(24):
        keep = _noise(side, dn, 32, 1 + (rep * 7) % 64) if rep % 3 else None       # every third repetition runs alone
        G = G0.clone(); lam = torch.zeros((B, S * N), device=DEV, dtype=dtype)
        dz = torch.zeros((B, nz), device=DEV, dtype=dtype)
        it = torch.zeros(B, dtype=torch.int32, device=DEV); ex = torch.zeros(B, dtype=torch.uint8, device=DEV)
        mg.kkt_solve_dz(S, C, N, G, Cm, g, cv, d["rho"], lam, dz, it, ex, cfg, batch=B)
        torch.cuda.synchronize()
        cur = (lam, dz, G, it, ex)
        if ref is None:
            ref = cur
            assert int(ex.sum()) == 0 and bool(torch.isfinite(lam).all())
        else:
            for a, b in zip(cur, ref):
                assert torch.equal(a, b), "repetition %d differs from the first run" % rep
        del keep
    # and the two-stage path (S, Pinv through HBM) gives the fused path's bits
    z = mg.sizes(S, C, N)
    G = G0.clone()
    Sg = torch.zeros((B, z["S"]), device=DEV, dtype=dtype); Pg = torch.zeros_like(Sg); gam = torch.zeros((B, z["gamma"]), device=DEV, dtype=dtype)
    lam = torch.zeros((B, S * N), device=DEV, dtype=dtype)
    it = torch.zeros(B, dtype=torch.int32, device=DEV); ex = torch.zeros(B, dtype=torch.uint8, device=DEV)
    mg.form_schur_system(S, C, N, G, Cm, g, cv, Sg, Pg, gam, d["rho"], batch=B)
    mg.pcg(S, N, Sg, Pg, gam, lam, it, ex, cfg, batch=B)
    torch.cuda.synchronize()
    assert torch.equal(lam, ref[0]) and torch.equal(G, ref[2]) and torch.equal(it, ref[3])


@pytest.mark.parametrize("N,world", [(128, 4), (256, 2), (512, 8)])
def test_knot_split_repeats_bit_identically(N, world):
    """The mailbox protocol under repetition: 30 back-to-back solves (alternating mailbox halves, growing epochs, no
    host synchronisation in between) must all return the first solve's bits."""
    d = synth.gen_kkt(N, 1, seed=321 + N)
    S32, P32, g32, _ = oracle.schur(S, C, N, d["G"][0], d["C"][0], d["g"][0], d["c"][0], d["rho"], np.float32)
    mail = torch.zeros((world, mg.mailbox_bytes() // 4), dtype=torch.int32, device=DEV)
    ptrs = [mail[r].data_ptr() for r in range(world)]
    Sg, Pg, gam = dev(S32), dev(P32), dev(g32)
    cfg = mg.PcgConfig(pcg_exit_tol=1e-5)
    outs, epoch = [], 0
    for rep in range(30):
        lam = torch.zeros(S * N, device=DEV)
        it = torch.zeros(world, dtype=torch.int32, device=DEV); ex = torch.zeros(world, dtype=torch.uint8, device=DEV)
        mg.knotsplit_pcg(S, N // world, world, 0, world, epoch, ptrs, Sg, Pg, gam, lam, it, ex, cfg)
        epoch += mg.epoch_stride(cfg)
        outs.append((lam, it))
    torch.cuda.synchronize()
    for lam, it in outs[1:]:
        assert torch.equal(lam, outs[0][0]) and torch.equal(it, outs[0][1])
    assert mg.knotsplit_last_timeout() is None
