"""Knot-range-split PCG on N GPUs (one process per GPU):
    python -m torch.distributed.run --nproc-per-node P --master-addr 127.0.0.1 tools/knotsplit_run.py [N] [reps]
Every rank forms the Schur rows of ITS knots from its extended input slice (no communication),
then all ranks run ONE persistent PCG kernel each; boundary rows and scalar sums cross NVLink as
tagged peer stores.  Rank 0 checks the gathered lambda against the CPU oracle and prints one JSON line."""
import json, os, sys
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mpcgpu_b200 as mg
from mpcgpu_b200 import synth

N = int(sys.argv[1]) if len(sys.argv) > 1 else 512
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 50
S, C = 14, 7
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
d = synth.gen_kkt(N, 1, seed=synth.BASE_SEED + N)
lo, hi = mg.knot_range(N, rank, world)
n = hi - lo
t1 = lambda a: torch.from_numpy(np.ascontiguousarray(a))
# my rows of S / Pinv / gamma from my slice of the inputs, no communication (ranges of up to 512 knots)
Sl, Pl, gl = mg.schur_rows_local(S, C, N, rank, world, t1(d["G"][0]), t1(d["C"][0]), t1(d["g"][0]), t1(d["c"][0]), d["rho"])
lam = torch.zeros(n * 14, device=dev)
it = torch.zeros(1, dtype=torch.int32, device=dev); ex = torch.zeros(1, dtype=torch.uint8, device=dev)
cfg = mg.PcgConfig(pcg_exit_tol=1e-4)
grp = mg.KnotSplitGroup()
for _ in range(3):
    lam.zero_(); grp.pcg(S, n, Sl, Pl, gl, lam, it, ex, cfg)
torch.cuda.synchronize(); dist.barrier()
# parity: gather lambda on rank 0
parts = [torch.empty_like(lam) for _ in range(world)]
dist.all_gather(parts, lam)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
e0.record()
for _ in range(reps):
    lam.zero_(); grp.pcg(S, n, Sl, Pl, gl, lam, it, ex, cfg)
e1.record(); torch.cuda.synchronize()
tt = torch.tensor([e0.elapsed_time(e1) / reps * 1e3], device=dev, dtype=torch.float64)
dist.all_reduce(tt, op=dist.ReduceOp.MAX)
if rank == 0:
    import oracle
    S32, P32, g32, _ = oracle.schur(S, C, N, d["G"][0], d["C"][0], d["g"][0], d["c"][0], d["rho"], np.float32)
    lam_s, _, _ = oracle.pcg(S, N, S32, P32, g32, np.zeros(S * N, np.float32), int(it.item()), 0.0)
    lam_o, it_o, _ = oracle.pcg(S, N, S32, P32, g32, np.zeros(S * N, np.float32), 10000, 1e-4)
    lam_g = torch.cat(parts).cpu().numpy()
    relerr = float(np.linalg.norm(lam_g - lam_s) / np.linalg.norm(lam_s))
    print(json.dumps({"metric": "knotsplit_pcg_solve_us", "knot_points": N, "n_gpus": world, "knots_per_gpu": n,
                      "value": float(tt[0]), "unit": "us (memset + solve, max over ranks, stream-timed)",
                      "iters": int(it.item()), "oracle_iters": it_o, "rel_err_vs_oracle_same_iters": relerr,
                      "us_per_iter": float(tt[0]) / max(int(it.item()), 1)}))
    assert abs(int(it.item()) - it_o) <= 1 and relerr < 5e-3
grp.close()
dist.destroy_process_group()
