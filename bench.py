#!/usr/bin/env python
"""bench.py -- headline benchmark of the KKT linear-system hot path (form_schur -> PCG).

Workload (BASELINE.json configs[4] shape, per GPU): `--batch` (4096) independent Kuka IIWA
64-knot KKT instances ("perturbed start states"), fp32, cold start lambda0 = 0, exit tol 1e-4,
rho = 1e-3.  A "step" = one fused KKT solve of the whole batch.  Weak scaling: every rank
owns its own 4096 instances, no data-path collective.

  value     whole-job solves/s with inputs resident in HBM (timed region: per step a D2D
            restore of G -- the API overwrites it with G^-1 -- a memset of lambda, and the
            fused kernel; CUDA events, barrier + synchronize on both sides, max over ranks)
  e2e       same metric through the host-buffer C-ABI solver (pinned host buffers, H2D + D2H
            inside the timed region)
  roofline  fused kernel alone: Def. B algorithmic bytes / CUDA-event duration
            against the measured HBM peak, plus achieved fp32 TFLOP/s (the kernel is
            FP32/latency bound, not HBM bound -- reported, not hidden)
  cpu_baseline  the reference's CPU work on the path = its QDLDL host solve (qdldl_solve_schur,
            include/qdldl/sqp.cuh:22-49: factor + solve on precomputed CSC values; the reference forms S
            on the GPU) on all host cores over the same 4096-instance batch, persistent thread pool;
            `with_schur_restatement` adds the oracle's CPU restatement of the Schur formation
  --impl reference   the reference arm: that QDLDL host solve with all host threads
  parity    sampled instances of the timed batch against the oracle / reference QDLDL (oracle/parity.py)
  warm_start, qdldl_single_core_us: SURVEY 8(d) legs, outside the timed region
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

S, C = 14, 7
F_SCHUR = 60e3          # flop / knot, reference algorithm
F_ITER = 2520.0         # flop / knot / PCG iteration


def def_b_bytes(N):
    return 3388 * N - 1596          # SURVEY 8(d) Def. B (fused solve, compulsory I/O)


def def_a_bytes(N):
    return 12908 * N - 1596         # Def. A (S, Pinv materialised)


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self._stop_evt, self.proc = index, [], threading.Event(), None

    def run(self):
        # ONE nvidia-smi process looping every 50 ms (a process per sample costs ~0.5 s each: a 40 ms timed
        # region would get a single sample)
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                if line.strip():
                    self.rows.append([x.strip() for x in line.split(",")])
                if self._stop_evt.is_set():
                    break
        except Exception:
            pass

    def stop(self):
        self._stop_evt.set()
        if self.proc is not None:
            try:
                self.proc.terminate()
            except Exception:
                pass
        self.join(timeout=6)
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows for i in range(4)
                          if len(r) >= 6 and r[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


def cpu_reference_run(N, B, steps, warmup, threads, keep=False):
    """The reference's CPU implementation of the path on `threads` host threads, on the bench batch itself.
    `value` leg: qdldl_solve_schur (QDLDL_factor + copy + QDLDL_solve, include/qdldl/sqp.cuh:22-49) of all B
    systems per step on precomputed CSC values / gamma -- the reference's QDLDL slot forms S on the GPU
    (include/qdldl/linsys_setup.cuh) and only factors / solves on the host, elimination tree once (:193).
    Threads, workspaces, pattern and tree persist across steps (oracle.CpuPool).  Second leg: the same preceded
    by the oracle's CPU restatement of that Schur formation (no preconditioner), on a bounded sample."""
    import oracle
    from mpcgpu_b200 import synth
    d = synth.gen_kkt_fast(N, B, seed=synth.BASE_SEED)
    pool = oracle.CpuPool(S, C, N, threads)
    val, gam = pool.schur_csc(d["G"], d["C"], d["g"], d["c"], d["rho"])        # preparation, not timed
    walls = []
    for it in range(warmup + steps):
        lam, wall = pool.qdldl(val, gam)
        if it >= warmup:
            walls.append(wall)
    wall = float(np.mean(walls))
    sample = min(B, max(threads * 16, 256))
    ws = []
    for it in range(3):
        _, w2, t_schur, t_qdldl = pool.kkt(d["G"][:sample], d["C"][:sample], d["g"][:sample], d["c"][:sample], d["rho"])
        if it >= 1:
            ws.append((w2, t_schur / (t_schur + t_qdldl)))
    pool.close()
    out = {"solves_per_s": B / wall, "ms_per_step": wall * 1e3, "batch": B,
           "with_schur": {"solves_per_s": sample / float(np.mean([w[0] for w in ws])), "sample": sample,
                          "restatement_share": float(np.mean([w[1] for w in ws]))}}
    if keep:
        out["lam_star"] = lam          # QDLDL (float) solutions of the whole batch: the warm-start leg perturbs them
    return out


def cpu_baseline_object(r, threads, B):
    return {"value": r["solves_per_s"], "unit": "solves/s", "cores": threads, "kind": "reference",
            "sample": "all %d instances of the bench batch per step; reference QDLDL factor + solve (oracle/_ref, the "
                      "reference's qdldl.c) on precomputed CSC values, persistent pool, elimination tree once; 100%% "
                      "reference-authored code in the timed region" % B,
            "qdldl_only": {"solves_per_s": r["solves_per_s"], "ms_per_step": r["ms_per_step"], "builder_code_share": 0.0},
            "with_schur_restatement": {"solves_per_s": r["with_schur"]["solves_per_s"], "sample": r["with_schur"]["sample"],
                                       "builder_code_share": r["with_schur"]["restatement_share"],
                                       "note": "adds the oracle's CPU restatement of form_schur_qdl_kernel (S, gamma, G^-1; "
                                               "no preconditioner) + CSC packing; the reference runs that stage on the GPU"}}


def qdldl_single_core(torch, dev):
    """SURVEY 8(d) CPU baseline (1) / BASELINE.json configs[0]: the reference's QDLDL host solve of ONE system on ONE
    core (median of 200 repetitions after 20 warm-ups) at N = 32 / 64 / 128 / 512, and the PCIe round trip its slot
    pays per solve (include/qdldl/sqp.cuh:268-273: blocking pageable D2H of nnz + s*N floats, H2D of s*N floats)."""
    import oracle
    from mpcgpu_b200 import synth
    out = {}
    for n in (32, 64, 128, 512):
        d = synth.gen_kkt(n, 1, seed=synth.BASE_SEED + n)
        S32, _, g32, _ = oracle.schur(S, C, n, d["G"][0], d["C"][0], d["g"][0], d["c"][0], d["rho"], np.float32)
        us, _ = oracle.qdldl_single_core_us(S, n, S32, g32, reps=200, warmup=20)
        nnz = (n - 1) * S * S + n * (S + 1) * S // 2
        rec = {"factor_solve_us": us, "nnz": nnz}
        if torch is not None:
            d_val = torch.zeros(nnz, device=dev); d_g = torch.zeros(S * n, device=dev); d_l = torch.zeros(S * n, device=dev)
            h_val = torch.zeros(nnz); h_g = torch.zeros(S * n); h_l = torch.zeros(S * n)      # pageable, like the reference's malloc
            ts = []
            for rep in range(220):
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                h_val.copy_(d_val); h_g.copy_(d_g); d_l.copy_(h_l)
                torch.cuda.synchronize()
                if rep >= 20:
                    ts.append((time.perf_counter() - t0) * 1e6)
            rec["pcie_round_trip_us"] = float(np.median(ts))
        out[str(n)] = rec
    return out


def _graph_time_us(torch, fn, reset, reps=20):
    """GPU-bound latency: a CUDA graph of `reps` (reset + fn) pairs replayed between two events on
    its own stream, so that no host launch overhead sits between the kernels."""
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        for _ in range(3):
            reset(); fn()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=st):
            for _ in range(reps):
                reset(); fn()
        gr.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st); gr.replay(); e1.record(st); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


def two_stage_batched(torch, mg, N, B, G0, Cm, g, cv, rho, tol, peak):
    """The two kernels of the reference's own structure (S, Pinv materialised in HBM, Def. A) on the
    same batch: each against the HBM roofline with its algorithmic bytes."""
    S, C = 14, 7
    dev = G0.device
    z = mg.sizes(S, C, N)
    G = G0.clone()
    Sg = torch.empty((B, z["S"]), device=dev); Pg = torch.empty_like(Sg)
    gam = torch.empty((B, z["gamma"]), device=dev)
    lam = torch.zeros((B, S * N), device=dev)
    it = torch.zeros(B, dtype=torch.int32, device=dev); ex = torch.ones(B, dtype=torch.uint8, device=dev)
    cfg = mg.PcgConfig(pcg_exit_tol=tol)
    ev = lambda: torch.cuda.Event(enable_timing=True)
    ts, tp = [], []
    for rep in range(6):
        G.copy_(G0); lam.zero_()
        a0, a1, b1 = ev(), ev(), ev()
        a0.record()
        mg.form_schur_system(S, C, N, G, Cm, g, cv, Sg, Pg, gam, rho, batch=B)
        a1.record()
        mg.pcg(S, N, Sg, Pg, gam, lam, it, ex, cfg, batch=B)
        b1.record()
        torch.cuda.synchronize()
        if rep >= 2:
            ts.append(a0.elapsed_time(a1)); tp.append(a1.elapsed_time(b1))
    ms_s, ms_p = float(np.mean(ts)), float(np.mean(tp))
    by_s, by_p = B * (8036 * N - 1596), B * 4872 * N
    return {"schur_kernel": {"ms": ms_s, "bytes": by_s, "GBps": by_s / ms_s / 1e6, "frac_hbm": by_s / ms_s / 1e6 / peak},
            "pcg_kernel": {"ms": ms_p, "bytes": by_p, "GBps": by_p / ms_p / 1e6, "frac_hbm": by_p / ms_p / 1e6 / peak,
                           "mean_iters": float(it.float().mean().item())},
            "two_stage_solves_per_s": B / ((ms_s + ms_p) * 1e-3)}


def single_solve_latency(torch, mg, N, tol):
    """Single-instance latency (BASELINE.json metric (i)), GPU-bound (CUDA-graph replay): PCG only
    (cluster kernel vs the reference-STYLE 4-grid-barrier kernel, our restatement of the absent
    GBD-PCG structure), Schur only, the fused solve, and the floor = same PCG launch with
    max_iter = 0 (launch + load of S / Pinv + first residual)."""
    from mpcgpu_b200 import synth
    dev = "cuda"
    d = synth.gen_kkt(N, 1, seed=synth.BASE_SEED + N)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    G0, Cm, g, cv = t(d["G"]), t(d["C"]), t(d["g"]), t(d["c"])
    z = mg.sizes(S, C, N)
    Sg = torch.empty((1, z["S"]), device=dev); Pg = torch.empty_like(Sg)
    gam = torch.empty((1, S * N), device=dev)
    G = G0.clone()
    mg.form_schur_system(S, C, N, G, Cm, g, cv, Sg, Pg, gam, d["rho"])
    lam = torch.zeros((1, S * N), device=dev)
    it = torch.zeros(1, dtype=torch.int32, device=dev); ex = torch.zeros(1, dtype=torch.uint8, device=dev)
    cfg = mg.PcgConfig(pcg_exit_tol=tol)
    cfg0 = mg.PcgConfig(pcg_exit_tol=tol, pcg_max_iter=0)
    ws = torch.empty(2 * S * N + 2 * N + 16, device=dev)
    out = {"N": N, "tol": tol, "timing": "CUDA-graph replay of 20 (memset + kernel) pairs, CUDA events"}
    out["pcg_us"] = _graph_time_us(torch, lambda: mg.pcg(S, N, Sg, Pg, gam, lam, it, ex, cfg), lambda: lam.zero_())
    out["pcg_iters"] = int(it.item())
    out["pcg_floor_us"] = _graph_time_us(torch, lambda: mg.pcg(S, N, Sg, Pg, gam, lam, it, ex, cfg0), lambda: lam.zero_())
    out["pcg_us_per_iter"] = (out["pcg_us"] - out["pcg_floor_us"]) / max(out["pcg_iters"], 1)
    try:
        out["pcg_refstyle_us"] = _graph_time_us(
            torch, lambda: mg.pcg(S, N, Sg, Pg, gam, lam, it, ex, cfg, refstyle=True, d_ws=ws), lambda: lam.zero_())
        out["pcg_refstyle_iters"] = int(it.item())
    except Exception as e:  # cooperative launches may refuse stream capture / co-residency
        out["pcg_refstyle_us"] = None
        out["pcg_refstyle_error"] = str(e)[:80]
    out["schur_us"] = _graph_time_us(torch, lambda: mg.form_schur_system(S, C, N, G, Cm, g, cv, Sg, Pg, gam, d["rho"]),
                                     lambda: G.copy_(G0))
    def reset_f():
        G.copy_(G0); lam.zero_()
    out["fused_us"] = _graph_time_us(torch, lambda: mg.kkt_solve(S, C, N, G, Cm, g, cv, d["rho"], lam, it, ex, cfg), reset_f)
    out["fused_iters"] = int(it.item())
    # host-visible latency of one fused solve: launch + kernel + blocking read-back of the iteration count
    # (what the reference reports as "linsys time": clock_gettime around launch + memcpys, include/pcg/sqp.cuh:269-286)
    for _ in range(20):
        reset_f(); mg.kkt_solve(S, C, N, G, Cm, g, cv, d["rho"], lam, it, ex, cfg); it.item()
    hv = []
    for _ in range(200):
        reset_f(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        mg.kkt_solve(S, C, N, G, Cm, g, cv, d["rho"], lam, it, ex, cfg)
        it.item()
        hv.append((time.perf_counter() - t0) * 1e6)
    out["fused_host_visible_us"] = float(np.median(hv))
    return out


def f64_leg(torch, mg, N, B, tol):
    """The fp64 twins (USE_DOUBLES 1): batched fused solves/s on B instances and one single-solve
    latency, same synthetic blocks as the fp32 workload converted to double."""
    from mpcgpu_b200 import synth
    dev = "cuda"
    d = synth.gen_kkt_fast(N, B, seed=synth.BASE_SEED)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(dev)
    G0, Cm, g, cv = t(d["G"]), t(d["C"]), t(d["g"]), t(d["c"])
    G = G0.clone()
    lam = torch.zeros((B, S * N), dtype=torch.float64, device=dev)
    it = torch.zeros(B, dtype=torch.int32, device=dev); ex = torch.ones(B, dtype=torch.uint8, device=dev)
    cfg = mg.PcgConfig(pcg_exit_tol=tol)
    ev = lambda: torch.cuda.Event(enable_timing=True)
    ts = []
    for rep in range(5):
        G.copy_(G0); lam.zero_()
        a0, a1 = ev(), ev()
        a0.record()
        mg.kkt_solve(S, C, N, G, Cm, g, cv, d["rho"], lam, it, ex, cfg, batch=B)
        a1.record(); torch.cuda.synchronize()
        if rep >= 2:
            ts.append(a0.elapsed_time(a1))
    ms = float(np.mean(ts))
    iters = it.float().mean().item()
    G1, l1 = G0[:1].clone(), lam[:1].clone()
    def reset():
        G1.copy_(G0[:1]); l1.zero_()
    us = _graph_time_us(torch, lambda: mg.kkt_solve(S, C, N, G1, Cm[:1], g[:1], cv[:1], d["rho"], l1, it[:1], ex[:1], cfg), reset)
    flops = N * (F_SCHUR * B + F_ITER * iters * B)
    return {"dtype": "f64", "knot_points": N, "batch": B, "exit_tol": tol, "kernel_ms": ms,
            "solves_per_s": B / (ms * 1e-3), "mean_iters": iters, "max_iter_exits": int(ex.sum().item()),
            "GBps_def_b": 2 * B * def_b_bytes(N) / (ms * 1e-3) / 1e9, "fp64_tflops": flops / (ms * 1e-3) / 1e12,
            "single_fused_us": us, "single_fused_iters": int(it[0].item())}


def knotsplit_bench(torch, mg, dist, rank, world, dev, N, tol, grp, reps=30, check=True):
    """BASELINE.json configs[3]: one N-knot horizon split by knot range over the `world` GPUs (NVLink peer
    mailboxes, no collective call inside the solve).  Stream-timed, max over ranks.  Rank 0 checks the gathered
    lambda against the oracle PCG on the full horizon (outside the timed region)."""
    from mpcgpu_b200 import synth
    d = synth.gen_kkt(N, 1, seed=synth.BASE_SEED + N)
    lo, hi = mg.knot_range(N, rank, world)
    n = hi - lo
    t1 = lambda a: torch.from_numpy(np.ascontiguousarray(a))
    Sl, Pl, gl = mg.schur_rows_local(S, C, N, rank, world, t1(d["G"][0]), t1(d["C"][0]), t1(d["g"][0]), t1(d["c"][0]), d["rho"])
    lam = torch.zeros(n * 14, device=dev)
    it = torch.zeros(1, dtype=torch.int32, device=dev); ex = torch.zeros(1, dtype=torch.uint8, device=dev)
    cfg = mg.PcgConfig(pcg_exit_tol=tol)
    for _ in range(3):
        lam.zero_(); grp.pcg(S, n, Sl, Pl, gl, lam, it, ex, cfg)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        lam.zero_(); grp.pcg(S, n, Sl, Pl, gl, lam, it, ex, cfg)
    e1.record(); torch.cuda.synchronize()
    tt = torch.tensor([e0.elapsed_time(e1) / reps * 1e3], device=dev, dtype=torch.float64)
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    iters = int(it.item())
    out = {"knot_points": N, "knots_per_gpu": n, "solve_us": float(tt[0]), "iters": iters,
           "us_per_iter": float(tt[0]) / max(iters, 1), "max_iter_exit": int(ex.item()),
           "exchange": "tagged 8-byte NVLink peer stores into per-GPU mailboxes, polled locally; no NCCL in the solve"}
    if check:
        def gather(x):
            parts = [torch.zeros_like(x) for _ in range(world)]
            dist.all_gather(parts, x)
            return torch.cat(parts).cpu().numpy() if rank == 0 else None
        lam_g, S_g, P_g, g_g = gather(lam), gather(Sl), gather(Pl), gather(gl)
        if rank == 0:
            import oracle
            l0 = np.zeros(S * N, np.float32)
            # stage-wise: the oracle PCG on the SAME S / Pinv / gamma rows the GPUs solved (their own Schur formation)
            _, it_o, _ = oracle.pcg(S, N, S_g, P_g, g_g, l0, 10000, tol)
            lam_s, _, _ = oracle.pcg(S, N, S_g, P_g, g_g, l0, iters, 0.0)
            floor = float(oracle.pcg_noise_floor(S, N, S_g, P_g, g_g, l0, iters))
            relv = float(np.linalg.norm(lam_g - lam_s) / np.linalg.norm(lam_s))
            # end to end (the oracle's own fp32 Schur formation: another rounding of S, amplified by kappa(S))
            S32, P32, g32, _ = oracle.schur(S, C, N, d["G"][0], d["C"][0], d["g"][0], d["c"][0], d["rho"], np.float32)
            lam_e, _, _ = oracle.pcg(S, N, S32, P32, g32, l0, iters, 0.0)
            out["parity"] = {"oracle_iters": it_o, "iters_within_1": bool(abs(iters - it_o) <= 1),
                             "lambda_rel_vs_oracle_same_rows_same_iters": relv, "fp32_floor": floor,
                             "within_max_1e-4_3x_floor": bool(relv <= max(1e-4, 3.0 * floor)),
                             "lambda_rel_vs_oracle_own_schur_rows": float(np.linalg.norm(lam_g - lam_e) / np.linalg.norm(lam_e)),
                             "schur_rows_rel_vs_oracle": float(np.linalg.norm(S_g - S32) / np.linalg.norm(S32))}
    return out


def long_horizon_one_gpu(torch, mg, dev, N, clusters, tol, reps=20):
    """Horizons beyond one cluster on ONE GPU: the knot-split kernel with all `clusters` knot ranges resident on this device
    (one cooperative launch of `clusters` clusters of 16 CTAs, 512 knots each; the mailboxes are local, the exchange goes
    through L2 instead of NVLink).  No single-cluster entry point takes N > 512.  Stream-timed; iteration count and lambda
    against the oracle PCG on the GPU's own S / Pinv / gamma rows."""
    import oracle
    from mpcgpu_b200 import synth
    d = synth.gen_kkt(N, 1, seed=synth.BASE_SEED + N)
    t1 = lambda a: torch.from_numpy(np.ascontiguousarray(a))
    rows = [mg.schur_rows_local(S, C, N, r, clusters, t1(d["G"][0]), t1(d["C"][0]), t1(d["g"][0]), t1(d["c"][0]), d["rho"]) for r in range(clusters)]
    Sg = torch.cat([r[0] for r in rows]); Pg = torch.cat([r[1] for r in rows]); gam = torch.cat([r[2] for r in rows])
    mail = torch.zeros((clusters, mg.mailbox_bytes() // 4), dtype=torch.int32, device=dev)
    ptrs = [mail[r].data_ptr() for r in range(clusters)]
    it = torch.zeros(clusters, dtype=torch.int32, device=dev); ex = torch.zeros(clusters, dtype=torch.uint8, device=dev)
    lam = torch.zeros(S * N, device=dev)
    cfg = mg.PcgConfig(pcg_exit_tol=tol)
    stride = mg.epoch_stride(cfg)
    epoch = 0

    def solve():
        nonlocal epoch
        lam.zero_()
        mg.knotsplit_pcg(S, N // clusters, clusters, 0, clusters, epoch, ptrs, Sg, Pg, gam, lam, it, ex, cfg)
        epoch += stride
    for _ in range(3):
        solve()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        solve()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / reps * 1e3
    iters = int(it[0].item())
    Sn, Pn, gn, l0 = Sg.cpu().numpy(), Pg.cpu().numpy(), gam.cpu().numpy(), np.zeros(S * N, np.float32)
    _, it_o, _ = oracle.pcg(S, N, Sn, Pn, gn, l0, 10000, tol)
    lam_s, _, _ = oracle.pcg(S, N, Sn, Pn, gn, l0, iters, 0.0)
    floor = float(oracle.pcg_noise_floor(S, N, Sn, Pn, gn, l0, iters))
    relv = float(np.linalg.norm(lam.cpu().numpy() - lam_s) / np.linalg.norm(lam_s))
    return {"knot_points": N, "clusters": clusters, "solve_us": us, "iters": iters, "us_per_iter": us / max(iters, 1),
            "max_iter_exit": int(ex.sum().item()), "oracle_iters": int(it_o), "lambda_rel_vs_oracle_same_iters": relv,
            "fp32_floor": floor, "within_max_1e-4_3x_floor": bool(relv <= max(1e-4, 3.0 * floor) and abs(iters - it_o) <= 1)}


def nccl_exchange_floor(torch, dist, rank, world, dev, iters=200):
    """SURVEY 8(e) "baseline to beat": what the knot split's two exchanges per PCG iteration cost when they go through
    NCCL -- all_reduce of the partial sums + send/recv of the 14-float boundary rows to both neighbours, twice per
    iteration, NO compute.  CUDA events, max over ranks; microseconds per PCG iteration."""
    red = torch.zeros(2, device=dev)
    out_l, in_l, out_r, in_r = (torch.zeros(14, device=dev) for _ in range(4))

    def one():
        for _ in range(2):
            dist.all_reduce(red)
            ops = []
            if rank > 0:
                ops += [dist.P2POp(dist.isend, out_l, rank - 1), dist.P2POp(dist.irecv, in_l, rank - 1)]
            if rank + 1 < world:
                ops += [dist.P2POp(dist.isend, out_r, rank + 1), dist.P2POp(dist.irecv, in_r, rank + 1)]
            for w in dist.batch_isend_irecv(ops):
                w.wait()
    for _ in range(20):
        one()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        one()
    e1.record(); torch.cuda.synchronize()
    tt = torch.tensor([e0.elapsed_time(e1) / iters * 1e3], device=dev, dtype=torch.float64)
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    return float(tt[0])


def knotsplit_leg(torch, mg, dist, rank, world, dev, tol):
    """Everything measured about the knot-range split at `world` GPUs; failures are reported, never fatal."""
    out = {}
    grp = None
    try:
        grp = mg.KnotSplitGroup()
        out["mailbox_one_way_us"] = grp.pingpong_us()
        out["note"] = ("exchange_floor = 2 one-way hops per PCG iteration (two exchanges); nccl_exchange = the same two "
                       "exchanges as torch.distributed NCCL calls (all_reduce + batched send/recv of the 14-float halos), "
                       "host-launched, no compute: the SURVEY 8(e) baseline")
        out["exchange_floor_us_per_iter"] = 2.0 * out["mailbox_one_way_us"]
        if 512 % world == 0 and 512 // world in (64, 128, 256):
            out["n512"] = knotsplit_bench(torch, mg, dist, rank, world, dev, 512, tol, grp)
        out["n512_per_gpu"] = knotsplit_bench(torch, mg, dist, rank, world, dev, 512 * world, tol, grp, reps=10)
    except Exception as e:      # noqa: BLE001
        out["error"] = repr(e)[:300]
    finally:
        if grp is not None:
            try:
                grp.close()
            except Exception:   # noqa: BLE001
                pass
    try:
        out["nccl_exchange_us_per_iter"] = nccl_exchange_floor(torch, dist, rank, world, dev)
    except Exception as e:      # noqa: BLE001
        out["nccl_error"] = repr(e)[:300]
    return out


def e2e_producer_leg(N, B, tol, steps, local):
    """VERDICT r1 "Next" #5: the same metric end to end with the KKT PRODUCER on the device -- the host ships each
    instance's iterate, start state and goal window (~7 KB) instead of its KKT blocks (149 KB) and reads back dz.
    Runs mpcgpu_b200/_plugins/e2e_producer (tools/e2e_producer.cu: reference GRiD dynamics in the batched producer ->
    fused Schur + PCG + dz), built where the reference's headers exist; one process per rank, all ranks at once."""
    exe = os.path.join(ROOT, "mpcgpu_b200", "_plugins", "e2e_producer")
    if not os.path.exists(exe):
        return {"unavailable": "mpcgpu_b200/_plugins/e2e_producer not built (needs the reference's dynamics headers)"}
    env = dict(os.environ, LD_LIBRARY_PATH="/usr/local/cuda/lib64:" + os.environ.get("LD_LIBRARY_PATH", ""))
    r = subprocess.run([exe, str(N), str(B), str(tol), str(steps), "3", str(local)], cwd=os.path.dirname(exe), env=env,
                       capture_output=True, text=True, timeout=600)
    if r.returncode != 0:
        return {"error": (r.stdout[-200:] + r.stderr[-200:])}
    return json.loads(r.stdout.strip().splitlines()[-1])


def parity_block(N, d, lam, it, ex, tol, samples):
    """Instances drawn from the TIMED batch itself (their lambda / iteration counts are those of the last timed step)
    against the oracle and the reference's QDLDL -- outside the timed region, the oracle as the checker only."""
    from oracle import parity
    B = lam.shape[0]
    idx = np.linspace(0, B - 1, min(samples, B)).astype(int)
    lam_h, it_h, ex_h = lam[idx].cpu().numpy(), it[idx].cpu().numpy(), ex[idx].cpu().numpy()
    rows = []
    for j, b in enumerate(idx):
        sysd = parity.systems(N, d["G"][b], d["C"][b], d["g"][b], d["c"][b], d["rho"])
        rows.append(parity.record(N, sysd, np.zeros(S * N, np.float32), tol, lam_h[j], int(it_h[j]), bool(ex_h[j])))
    out = parity.summarize(rows)
    out["instances"] = [int(b) for b in idx]
    return out


def variant2_leg():
    """Informational: the same workload with MPCGPU_B200_PCG_VARIANT=2 (read once per process, hence a child process:
    tools/quick_fused.py).  Not the default and not the headline: DESIGN.md 3 has the accuracy it costs."""
    import subprocess
    env = dict(os.environ, MPCGPU_B200_PCG_VARIANT="2")
    try:
        r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "quick_fused.py"), "--child", "variant2"], env=env,
                           capture_output=True, text=True, timeout=300)
        out = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    except Exception as e:      # informational leg: never fails the bench line
        return {"error": str(e)[:200]}
    return {"kernel_ms": out["kernel_ms"], "solves_per_s": out["solves_per_s"], "iters_sum": out["iters_sum"],
            "note": "opt-in (MPCGPU_B200_PCG_VARIANT=2): S p by recurrence, r.rt reduction hidden behind the matvec; meets the operating-point "
                    "criteria but is 5-20x farther from the exact-arithmetic iterate than the default at tight tolerances"}


def warm_start_leg(torch, mg, N, B, G, G0, Cm, g, cv, rho, lam_star, it, ex, tol, steps):
    """the reference's operating point is a warm-started lambda (include/mpcsim.cuh:227,383).  Same batch,
    lambda0 = lambda* (1 + N(0, 1e-2)) with lambda* = the reference QDLDL solutions; kernel time by CUDA events."""
    rng = np.random.default_rng(12345)
    warm = (lam_star * (1.0 + 1e-2 * rng.standard_normal(lam_star.shape))).astype(np.float32)
    lam0 = torch.from_numpy(warm).to(G.device)
    lam = lam0.clone()
    cfg = mg.PcgConfig(pcg_exit_tol=tol)
    ts = []
    for k in range(3 + steps):
        G.copy_(G0); lam.copy_(lam0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        mg.kkt_solve(S, C, N, G, Cm, g, cv, rho, lam, it, ex, cfg, batch=B)
        e1.record(); torch.cuda.synchronize()
        if k >= 3:
            ts.append(e0.elapsed_time(e1))
    ms = float(np.mean(ts))
    iters = it.float()
    return {"solves_per_s": B / (ms * 1e-3), "kernel_ms": ms, "mean_iters": float(iters.mean().item()),
            "max_iters": int(iters.max().item()), "max_iter_exits": int(ex.sum().item()),
            "lambda0": "lambda*(1 + N(0, 1e-2)), lambda* = reference QDLDL (float) on the oracle's S"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--knots", type=int, default=64)
    ap.add_argument("--tol", type=float, default=1e-4)
    ap.add_argument("--no-extras", action="store_true")
    ap.add_argument("--parity-samples", type=int, default=32)
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    N, B = a.knots, a.batch
    config = {"workload": "%d independent Kuka IIWA (14-state, 7-control) %d-knot KKT instances per GPU, "
                          "fused Schur formation + PCG (SS preconditioner), cold start, exit tol %g, rho 1e-3"
                          % (B, N, a.tol),
              "knot_points": N, "batch_per_gpu": B, "exit_tol": a.tol,
              "cache": "inputs (%.2f GB per GPU) larger than L2" % (B * def_b_bytes(N) / 1e9)}

    if a.impl == "reference":
        if rank != 0:
            return
        threads = os.cpu_count() or 1
        r = cpu_reference_run(N, B, a.steps, max(a.warmup, 1), threads)
        line = {"impl": "reference", "metric": "batched_kkt_solves_per_sec", "value": r["solves_per_s"],
                "unit": "solves/s", "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
                "cpu_baseline": cpu_baseline_object(r, threads, B),
                "e2e": {"value": r["solves_per_s"], "unit": "solves/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "note": "the reference's CPU work on this path is its QDLDL host solve (direct LDL^T, no iteration); its "
                        "Schur formation runs on the GPU and is NOT in this figure -- see cpu_baseline.with_schur_restatement"}
        print(json.dumps(line))
        return

    import torch
    import mpcgpu_b200 as mg
    from mpcgpu_b200 import synth
    assert torch.cuda.is_available(), "bench.py needs a GPU (no CPU fallback)"
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)

    # every rank: its own perturbed instances of the SAME base trajectory (configs[4]: one problem, perturbed
    # start states), so that the per-GPU work is the same at every N (weak scaling); rank 0's are those of N = 1
    d = synth.gen_kkt_fast(N, B, seed=synth.BASE_SEED, perturb_seed=(rank if rank else None))
    pin = lambda x: torch.from_numpy(np.ascontiguousarray(x)).pin_memory()
    hG, hC, hg, hc = pin(d["G"]), pin(d["C"]), pin(d["g"]), pin(d["c"])
    G0, Cm, g, cv = hG.to(dev), hC.to(dev), hg.to(dev), hc.to(dev)
    G = G0.clone()
    lam = torch.zeros((B, S * N), device=dev)
    it = torch.zeros(B, dtype=torch.int32, device=dev); ex = torch.ones(B, dtype=torch.uint8, device=dev)
    cfg = mg.PcgConfig(pcg_exit_tol=a.tol)
    rho = d["rho"]

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def step(k0=None, k1=None):
        G.copy_(G0); lam.zero_()
        if k0 is not None: k0.record()
        mg.kkt_solve(S, C, N, G, Cm, g, cv, rho, lam, it, ex, cfg, batch=B)
        if k1 is not None: k1.record()

    for _ in range(max(a.warmup, 3)):
        step()
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler: sampler.start()
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n0 = mg.launch_count()
    barrier()
    e0.record()
    for k in range(a.steps):
        step(*kev[k])
    e1.record()
    barrier()
    launches = mg.launch_count() - n0
    total_ms = e0.elapsed_time(e1)
    kern_ms = float(np.mean([x.elapsed_time(y) for x, y in kev]))
    iters = it.cpu().numpy().astype(np.float64)
    per_rank = None
    if dist is not None:
        # every rank solves its OWN perturbed instances: the step time is the max over ranks
        mine = torch.tensor([kern_ms, float(iters.mean())], device=dev, dtype=torch.float64)
        allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        per_rank = {"kernel_ms": [round(float(x[0]), 4) for x in allr], "mean_pcg_iters": [round(float(x[1]), 2) for x in allr]}
        tt = torch.tensor([total_ms, kern_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        total_ms, kern_ms = float(tt[0]), float(tt[1])
    assert int(ex.sum().item()) == 0, "PCG hit max_iter in the bench workload"

    # ---- e2e through the host-buffer C-ABI solver ------------------------------------------
    hs = mg.HostSolver(S, C, N, B, device=local)
    hl = torch.zeros((B, S * N)).pin_memory()
    hl0 = torch.zeros((B, S * N)).pin_memory()      # warm start of every step (copied in, never written)
    hi = torch.zeros(B, dtype=torch.int32).pin_memory(); he = torch.ones(B, dtype=torch.uint8).pin_memory()
    for _ in range(2):
        hs.solve(hG, hC, hg, hc, rho, hl, hi, he, cfg, B, h_lambda0=hl0)
    barrier()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        hs.solve(hG, hC, hg, hc, rho, hl, hi, he, cfg, B, h_lambda0=hl0)
    torch.cuda.synchronize()
    e2e_s = (time.perf_counter() - t0) / a.steps
    if dist is not None:
        tt = torch.tensor([e2e_s], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e_s = float(tt[0])
    clocks = sampler.stop() if sampler else None
    assert torch.equal(hl.to(dev), lam), "host path and device path disagree"
    barrier()
    ep = e2e_producer_leg(N, B, a.tol, a.steps, local)          # every rank at once (shared host), own process per rank
    if dist is not None and "solves_per_s" in ep:
        tt = torch.tensor([ep["solves_per_s"]], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MIN)
        ep["solves_per_s_all_gpus"] = float(tt[0]) * world        # slowest rank x ranks
    elif "solves_per_s" in ep:
        ep["solves_per_s_all_gpus"] = ep["solves_per_s"]
    h2d = 4 * B * (hG[0].numel() + hC[0].numel() + hg[0].numel() + hc[0].numel() + S * N)
    d2h = B * (4 * S * N + 4 + 1)
    hs.close()
    # what bounds e2e: the host->device copy of the KKT blocks.  Measured pinned H2D rate of this box, same process:
    # rank 0 alone, then ALL ranks copying at the same time (they share the host's PCIe root / memory).
    probe_h = torch.empty(256 << 20, dtype=torch.uint8).pin_memory()
    probe_d = torch.empty_like(probe_h, device=dev)

    def h2d_rate():
        pe0, pe1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        probe_d.copy_(probe_h, non_blocking=True); torch.cuda.synchronize()
        pe0.record()
        for _ in range(4):
            probe_d.copy_(probe_h, non_blocking=True)
        pe1.record(); torch.cuda.synchronize()
        return 4 * probe_h.numel() / (pe0.elapsed_time(pe1) * 1e-3) / 1e9

    h2d_gbps = h2d_rate() if rank == 0 else 0.0
    h2d_conc = h2d_gbps
    if dist is not None:
        barrier()
        mine = h2d_rate()                       # every rank at once
        tt = torch.tensor([mine], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MIN)
        h2d_conc = float(tt[0])
    del probe_h, probe_d

    ks = None
    if world > 1 and not a.no_extras:
        ks = knotsplit_leg(torch, mg, dist, rank, world, dev, a.tol)
    if rank == 0:
        peak, how = load_peaks()
        ms_per_step = total_ms / a.steps
        ach = B * def_b_bytes(N) / (kern_ms * 1e-3) / 1e9
        flops = N * (F_SCHUR * B + F_ITER * iters.sum())
        props = torch.cuda.get_device_properties(local)
        mhz = (clocks or {}).get("sm_mhz") or (clocks or {}).get("sm_max_mhz") or 1965.0
        fp32_peak = props.multi_processor_count * 128 * 2 * mhz * 1e6 / 1e12      # FMA lanes x 2 flop x SM clock under load
        prof = os.path.join(ROOT, "profiles", "traffic.json")
        tj = json.load(open(prof)) if os.path.exists(prof) else {}
        traffic = tj.get("fused_kernel_dram_bytes_per_launch")
        ffma = tj.get("fused_kernel_fp32_flop_executed_per_launch")     # from ncu sass_thread_inst_executed (same workload)
        line = {"metric": "batched_kkt_solves_per_sec", "value": world * B / (ms_per_step * 1e-3), "unit": "solves/s",
                "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": ms_per_step,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic", "config": config, "clocks": clocks, "gpu_launches": int(launches),
                "pcg_iters": {"mean": float(iters.mean()), "min": int(iters.min()), "max": int(iters.max())},
                "e2e": {"value": world * B / e2e_s, "unit": "solves/s", "h2d_bytes_per_step": int(h2d),
                        "d2h_bytes_per_step": int(d2h), "bound": "pcie_h2d", "h2d_achieved_GBps": h2d / e2e_s / 1e9,
                        "h2d_peak_GBps": h2d_conc, "h2d_frac": h2d / e2e_s / 1e9 / h2d_conc, "h2d_peak_GBps_rank0_alone": h2d_gbps,
                        "note": "per-rank figures; peak = pinned 256 MiB H2D copies timed in this process with ALL ranks "
                                "copying at the same time (min over ranks): the ranks share the host's PCIe / memory"},
                "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                             "traffic": traffic, "traffic_source": "static: " + tj.get("source", "profiles/traffic.json") if traffic else None,
                             "peak_source": how + " (MEASURED_PEAKS.json hbm_gbs)" if how == "measured" else how,
                             "kernel": "fused_kernel (schur_rows + pcg_rows)", "kernel_ms": kern_ms,
                             "bytes_def": "Def. B 3388N-1596 B/instance (SURVEY 8d)",
                             "def_a_equiv_GBps": B * def_a_bytes(N) / (kern_ms * 1e-3) / 1e9,
                             "fp32_tflops": flops / (kern_ms * 1e-3) / 1e12,
                             "fp32_peak_tflops": fp32_peak, "fp32_frac": flops / (kern_ms * 1e-3) / 1e12 / fp32_peak,
                             "fp32_tflops_executed": (ffma / (kern_ms * 1e-3) / 1e12) if ffma else None,
                             "fp32_frac_executed": (ffma / (kern_ms * 1e-3) / 1e12 / fp32_peak) if ffma else None,
                             "note": "FP32/latency bound by design: arithmetic intensity >= 10 flop/B at 0 iterations (81 flop/B at the "
                                     "bench's 84 iterations, ridge 11 flop/B): at 100% FP32 issue the HBM fraction could reach ~14%, so "
                                     "north_star's 60% HBM target does not apply to a fused kernel; fp32_frac = nominal flops "
                                     "(60 kflop/knot Schur + 2520/knot/iteration) / (SMs x 128 FMA lanes x 2 x measured SM clock); "
                                     "fp32_frac_executed = the kernel's own executed FFMA/FADD/FMUL thread instructions (ncu, static "
                                     "capture of the same workload) over the same peak"}}
        if world == 1:
            threads = os.cpu_count() or 1
            r = cpu_reference_run(N, B, 3, 1, threads, keep=True)
            line["cpu_baseline"] = cpu_baseline_object(r, threads, B)
            line["parity"] = parity_block(N, d, lam, it, ex, a.tol, a.parity_samples)
            line["warm_start"] = warm_start_leg(torch, mg, N, B, G, G0, Cm, g, cv, rho, r["lam_star"], it, ex, a.tol, a.steps)
            if not a.no_extras:
                # informational legs: a failure in one of them is recorded in its key and never costs the bench line
                def leg(name, fn):
                    try:
                        line[name] = fn()
                    except Exception as e:
                        line[name] = {"error": "%s: %s" % (type(e).__name__, str(e)[:300])}
                leg("qdldl_single_core_us", lambda: qdldl_single_core(torch, dev))
                leg("two_stage", lambda: two_stage_batched(torch, mg, N, B, G0, Cm, g, cv, rho, a.tol, peak))
                leg("single_solve", lambda: [single_solve_latency(torch, mg, n, a.tol) for n in (64, 128, 512)])
                leg("f64", lambda: f64_leg(torch, mg, N, min(B, 1024), a.tol))
                leg("pcg_variant2_opt_in", variant2_leg)
                leg("long_horizon_one_gpu", lambda: [long_horizon_one_gpu(torch, mg, dev, n, c, a.tol) for n, c in ((1024, 2), (2048, 4))])
        line["e2e_producer"] = ep
        if per_rank is not None:
            line["per_rank"] = per_rank
        if ks is not None:
            line["knot_split"] = ks
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
