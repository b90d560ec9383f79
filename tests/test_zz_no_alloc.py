"""GPU test (-m gpu) of the allocation-free entry point mpcgpu_b200_pcg_refstyle_noalloc_f32 (no counterpart in the
reference): called through the C ABI (ctypes), its lambda / iteration
count / flag must equal those of the existing entry point mpcgpu_b200_pcg_refstyle_f32 (caller-owned d_ws) on the same
system -- on calls in a row that share the library's static scratch (different horizons, so a call finds the
previous call's contents there).  Named test_zz_* so that pytest collects it last."""
import ctypes

import numpy as np
import pytest
import torch

import mpcgpu_b200 as mg
from mpcgpu_b200 import synth
from helpers import S, C, rel

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _system(N, seed):
    d = synth.gen_kkt(N, 1, seed=seed)
    z = mg.sizes(S, C, N)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(DEV)
    G, Cm, g, cv = t(d["G"]), t(d["C"]), t(d["g"]), t(d["c"])
    Sg = torch.zeros((1, z["S"]), device=DEV); Pg = torch.zeros_like(Sg); gam = torch.zeros((1, z["gamma"]), device=DEV)
    mg.form_schur_system(S, C, N, G, Cm, g, cv, Sg, Pg, gam, d["rho"])
    torch.cuda.synchronize()
    return Sg, Pg, gam


def _noalloc(N, Sg, Pg, gam, lam, it, ex, max_iter, tol):
    L = mg.lib()
    p = lambda x: ctypes.c_void_p(x.data_ptr())
    rc = L.mpcgpu_b200_pcg_refstyle_noalloc_f32(ctypes.c_uint32(S), ctypes.c_uint32(N), p(Sg), p(Pg), p(gam), p(lam),
                                                ctypes.c_uint32(max_iter), ctypes.c_float(tol), p(it), p(ex),
                                                ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    return rc


def test_noalloc_entry_point_matches_the_existing_one_on_calls_in_a_row():
    L = mg.lib()
    L.mpcgpu_b200_noalloc_max_knots.restype = ctypes.c_uint32
    assert L.mpcgpu_b200_noalloc_max_knots() == 512
    tol, max_iter = 1e-5, 10000
    cfg = mg.PcgConfig(pcg_exit_tol=tol, pcg_max_iter=max_iter)
    # outputs of the new entry point are allocated up front: nothing is allocated around its calls
    cases = [(128, 11), (64, 12), (128, 13), (512, 14)]
    systems = [_system(N, seed) for N, seed in cases]
    outs = [(torch.zeros((1, S * N), device=DEV), torch.full((1,), -1, dtype=torch.int32, device=DEV),
             torch.full((1,), 7, dtype=torch.uint8, device=DEV)) for N, _ in cases]
    for (N, _), (Sg, Pg, gam), (lam, it, ex) in zip(cases, systems, outs):          # calls in a row, one stream
        assert _noalloc(N, Sg, Pg, gam, lam, it, ex, max_iter, tol) == 0, L.mpcgpu_b200_last_error()
    torch.cuda.synchronize()
    for (N, _), (Sg, Pg, gam), (lam, it, ex) in zip(cases, systems, outs):
        lam_ref = torch.zeros((1, S * N), device=DEV)
        it_ref = torch.full((1,), -1, dtype=torch.int32, device=DEV); ex_ref = torch.full((1,), 7, dtype=torch.uint8, device=DEV)
        mg.pcg(S, N, Sg, Pg, gam, lam_ref, it_ref, ex_ref, cfg, refstyle=True)           # existing entry point, own d_ws
        torch.cuda.synchronize()
        assert int(it[0]) == int(it_ref[0]) and int(it[0]) >= 1 and int(ex[0]) == int(ex_ref[0]) == 0
        assert rel(lam.cpu().numpy(), lam_ref.cpu().numpy()) < 1e-4            # the tolerance of the other PCG tests
        assert torch.equal(lam, lam_ref)                                       # same kernel body, same order: same bits


def test_noalloc_entry_point_rejects_sizes_beyond_its_static_scratch():
    L = mg.lib()
    L.mpcgpu_b200_last_error.restype = ctypes.c_char_p
    Sg, Pg, gam = _system(8, 3)
    lam = torch.zeros((1, S * 8), device=DEV)
    it = torch.zeros(1, dtype=torch.int32, device=DEV); ex = torch.zeros(1, dtype=torch.uint8, device=DEV)
    assert _noalloc(513, Sg, Pg, gam, lam, it, ex, 10, 1e-4) == 1
    assert b"at most" in L.mpcgpu_b200_last_error() and b"512" in L.mpcgpu_b200_last_error()
    rc = L.mpcgpu_b200_pcg_refstyle_noalloc_f32(ctypes.c_uint32(12), ctypes.c_uint32(8), None, None, None, None,
                                                ctypes.c_uint32(1), ctypes.c_float(1e-4), None, None, None)
    assert rc == 1 and b"state_size must be 14" in L.mpcgpu_b200_last_error()
    assert _noalloc(8, Sg, Pg, gam, lam, it, ex, 100, 1e-4) == 0               # and still works afterwards
    torch.cuda.synchronize()
    assert int(it[0]) >= 1 and bool(torch.isfinite(lam).all())
