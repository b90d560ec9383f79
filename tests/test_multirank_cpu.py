"""CPU tests (world_size 2, gloo) of the multi-rank HOST logic: instance slabs of the batched
mode (no data-path collective), knot ranges / extended Schur slices of the knot-split mode, and
the rule that every rank's extended slice reproduces the full-horizon Schur rows (checked with
the CPU oracle standing in for the device kernels -- the oracle is the checker here, the
product's kernels are covered by the -m gpu tests)."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def _worker(rank, world, port, N):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle
    import mpcgpu_b200 as mg
    from mpcgpu_b200 import synth
    S, C = 14, 7
    d = synth.gen_kkt(N, 1, seed=123)
    lo, hi = mg.knot_range(N, rank, world)
    lo_e, hi_e, off = mg.schur_slice(N, rank, world)
    ne = hi_e - lo_e
    ze = mg.sizes(S, C, ne)
    sl = lambda a, w, key: a[0][lo_e * w: lo_e * w + ze[key]]
    Ss, Ps, gs, _ = oracle.schur(S, C, ne, sl(d["G"], 245, "G"), sl(d["C"], 294, "C"), sl(d["g"], 21, "g"),
                                 sl(d["c"], 14, "c"), d["rho"], np.float32)
    n = hi - lo
    mine = torch.from_numpy(np.concatenate([Ss[off * 588:(off + n) * 588], Ps[off * 588:(off + n) * 588],
                                            gs[off * 14:(off + n) * 14]]))
    parts = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine)
    # max-over-ranks timing reduction used by bench.py
    t = torch.tensor([float(rank + 1)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    assert float(t[0]) == world
    if rank == 0:
        Sf, Pf, gf, _ = oracle.schur(S, C, N, d["G"][0], d["C"][0], d["g"][0], d["c"][0], d["rho"], np.float32)
        for r in range(world):
            a, b = mg.knot_range(N, r, world)
            p = parts[r].numpy()
            m = b - a
            assert np.array_equal(p[:m * 588], Sf[a * 588:b * 588])
            assert np.array_equal(p[m * 588:2 * m * 588], Pf[a * 588:b * 588])
            assert np.array_equal(p[2 * m * 588:], gf[a * 14:b * 14])
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("N", [32, 64])
def test_knot_ranges_and_schur_slices_world2(N):
    port = 29600 + (os.getpid() % 200) + N
    mp.spawn(_worker, args=(2, port, N), nprocs=2, join=True)


def test_knot_range_rules():
    import mpcgpu_b200 as mg
    assert mg.knot_range(512, 3, 4) == (384, 512)
    assert mg.schur_slice(512, 0, 4) == (0, 129, 0)
    assert mg.schur_slice(512, 3, 4) == (382, 512, 2)
    with pytest.raises(mg.MpcGpuError):
        mg.knot_range(100, 0, 3)
    with pytest.raises(mg.MpcGpuError):
        mg.knot_range(24, 0, 2)          # 12 knots per rank: not a supported tile
    assert mg.epoch_stride(mg.PcgConfig(pcg_max_iter=100)) > 2 * 100 + 8


def test_rank_shards_share_the_base_trajectory():
    """bench.py's weak-scaling shards: every rank holds different perturbed instances of the SAME base trajectory
    (instance 0 of a shard is the unperturbed base), reproducibly."""
    import numpy as np
    from mpcgpu_b200 import synth
    r0 = synth.gen_kkt_fast(8, 6)
    r1 = synth.gen_kkt_fast(8, 6, perturb_seed=1)
    r1b = synth.gen_kkt_fast(8, 6, perturb_seed=1)
    r2 = synth.gen_kkt_fast(8, 6, perturb_seed=2)
    for k in ("G", "C", "g", "c"):
        assert np.array_equal(r0[k][0], r1[k][0]) and np.array_equal(r1[k][0], r2[k][0])     # the base
        assert np.array_equal(r1[k], r1b[k])
    assert not np.array_equal(r1["C"][1:], r2["C"][1:]) and not np.array_equal(r0["C"][1:], r1["C"][1:])
