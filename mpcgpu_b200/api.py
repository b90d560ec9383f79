"""Host-side mirror of the reference's linear-solver interface, over the C ABI
(``include/mpcgpu_b200.h`` -> ``libmpcgpu_b200.so``).

Same names, argument order and meaning as the reference (torch CUDA tensors stand in for the
reference's raw device pointers; torch is plumbing for device memory / streams only):

* :func:`form_schur_system` -- ``include/pcg/linsys_setup.cuh:620-657``
* :func:`pcg`               -- launch at ``include/pcg/sqp.cuh:275`` + read-back ``:276-277``
* :func:`kkt_solve`         -- the two fused (``include/pcg/sqp.cuh:243-277`` as one launch)
* :class:`PcgConfig`        -- ``pcg_config<T>`` (``include/mpcsim.cuh:250-253``)
* :class:`HostSolver`       -- solver object for KKT blocks that live in host memory

There is NO CPU fallback: if the CUDA library is missing or no GPU is present, every
compute entry point raises.
"""
import ctypes
import os
from dataclasses import dataclass

import torch

from . import build as _build

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class MpcGpuError(RuntimeError):
    pass


def lib():
    """Load (once) the in-tree CUDA library; raises loudly if it is absent."""
    global _LIB
    if _LIB is None:
        path = os.environ.get("MPCGPU_B200_LIB", _build.LIB)   # an alternative BUILD of this library (A/B experiments)
        if not os.path.exists(path):
            raise MpcGpuError(
                "%s not found: build it with `python -m mpcgpu_b200.build` "
                "(there is no CPU fallback)" % path)
        L = ctypes.CDLL(path)
        L.mpcgpu_b200_last_error.restype = ctypes.c_char_p
        L.mpcgpu_b200_workspace_bytes.restype = ctypes.c_size_t
        L.mpcgpu_b200_launch_count.restype = ctypes.c_uint64
        if L.mpcgpu_b200_abi_version() != 1:
            raise MpcGpuError("ABI version mismatch")
        _LIB = L
    return _LIB


def _check(rc, what):
    if rc != 0:
        raise MpcGpuError("%s failed (cudaError %d): %s"
                          % (what, rc, lib().mpcgpu_b200_last_error().decode()))


def _ptr(t, dtype, what, numel=None):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise MpcGpuError("%s must be a CUDA tensor (no CPU fallback)" % what)
    if t.dtype != dtype or not t.is_contiguous():
        raise MpcGpuError("%s must be contiguous %s" % (what, dtype))
    if numel is not None and t.numel() < numel:
        raise MpcGpuError("%s too small: %d < %d elements" % (what, t.numel(), numel))
    return ctypes.c_void_p(t.data_ptr())


def _prec(t, what):
    """linsys_t of a call = dtype of its matrices: float32 (settings.cuh USE_DOUBLES 0) -> the _f32 entry
    points, float64 (USE_DOUBLES 1) -> the _f64 twins.  Returns (dtype, suffix, ctypes scalar type)."""
    if isinstance(t, torch.Tensor) and t.dtype == torch.float64:
        return torch.float64, "f64", ctypes.c_double
    if isinstance(t, torch.Tensor) and t.dtype == torch.float32:
        return torch.float32, "f32", ctypes.c_float
    raise MpcGpuError("%s must be a float32 or float64 CUDA tensor" % what)


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def sizes(s, c, N):
    """Element counts of the reference buffers (include/pcg/sqp.cuh:54-58,111,130)."""
    return {
        "G": (s * s + c * c) * N - c * c, "C": (s * s + s * c) * (N - 1),
        "g": (s + c) * N - c, "c": s * N, "S": 3 * s * s * N, "Pinv": 3 * s * s * N,
        "gamma": s * N, "lambda": s * N,
    }


@dataclass
class PcgConfig:
    """pcg_config<T> (include/mpcsim.cuh:250-253)."""
    pcg_block: int = 128
    pcg_exit_tol: float = 1e-4
    pcg_max_iter: int = 10000   # include/common/settings.cuh:133 (this fork)


def form_schur_system(state_size, control_size, knot_points, d_G_dense, d_C_dense, d_g, d_c,
                      d_S, d_Pinv, d_gamma, rho, batch=1):
    """form_schur_system<T>: writes d_S, d_Pinv, d_gamma; d_G_dense becomes G^-1.
    T = float / double by the dtype of d_G_dense."""
    s, c, N = state_size, control_size, knot_points
    z = sizes(s, c, N)
    f, sfx, cT = _prec(d_G_dense, "d_G_dense")
    rc = getattr(lib(), "mpcgpu_b200_schur_" + sfx)(
        ctypes.c_uint32(s), ctypes.c_uint32(c), ctypes.c_uint32(N), ctypes.c_uint32(batch),
        _ptr(d_G_dense, f, "d_G_dense", z["G"] * batch),
        _ptr(d_C_dense, f, "d_C_dense", z["C"] * batch) if N > 1 else None,
        _ptr(d_g, f, "d_g", z["g"] * batch), _ptr(d_c, f, "d_c", z["c"] * batch),
        cT(rho), _ptr(d_S, f, "d_S", z["S"] * batch),
        _ptr(d_Pinv, f, "d_Pinv", z["S"] * batch), _ptr(d_gamma, f, "d_gamma", z["gamma"] * batch),
        _stream())
    _check(rc, "form_schur_system")


def pcg(state_size, knot_points, d_S, d_Pinv, d_gamma, d_lambda, d_pcg_iters, d_pcg_exit,
        config: PcgConfig, batch=1, refstyle=False, d_ws=None):
    """pcg<float,STATE_SIZE,KNOT_POINTS>: d_lambda in (warm start) / out.
    d_pcg_iters: uint32 (int32 tensor storage), d_pcg_exit: uint8/bool tensor."""
    s, N = state_size, knot_points
    z = sizes(s, 7, N)
    f, sfx, cT = _prec(d_S, "d_S")
    if d_pcg_iters.dtype != torch.int32 or d_pcg_exit.dtype not in (torch.uint8, torch.bool):
        raise MpcGpuError("d_pcg_iters must be int32 and d_pcg_exit uint8/bool")
    args = [_ptr(d_S, f, "d_S", z["S"] * batch), _ptr(d_Pinv, f, "d_Pinv", z["S"] * batch),
            _ptr(d_gamma, f, "d_gamma", z["gamma"] * batch),
            _ptr(d_lambda, f, "d_lambda", z["lambda"] * batch),
            ctypes.c_uint32(config.pcg_max_iter), cT(config.pcg_exit_tol),
            _ptr(d_pcg_iters, torch.int32, "d_pcg_iters", batch),
            _ptr(d_pcg_exit, d_pcg_exit.dtype, "d_pcg_exit", batch)]
    if refstyle:
        if batch != 1 or sfx != "f32":
            raise MpcGpuError("reference-style kernel is single-instance, fp32")
        need = lib().mpcgpu_b200_workspace_bytes(ctypes.c_uint32(s), ctypes.c_uint32(7),
                                                 ctypes.c_uint32(N), ctypes.c_uint32(1))
        if d_ws is None:
            d_ws = torch.empty(need // 4, dtype=f, device=d_S.device)
        rc = lib().mpcgpu_b200_pcg_refstyle_f32(
            ctypes.c_uint32(s), ctypes.c_uint32(N), *args, _ptr(d_ws, f, "d_ws", need // 4), _stream())
    else:
        rc = getattr(lib(), "mpcgpu_b200_pcg_" + sfx)(ctypes.c_uint32(s), ctypes.c_uint32(N),
                                                      ctypes.c_uint32(batch), *args, None, _stream())
    _check(rc, "pcg")


def kkt_solve(state_size, control_size, knot_points, d_G_dense, d_C_dense, d_g, d_c, rho,
              d_lambda, d_pcg_iters, d_pcg_exit, config: PcgConfig, batch=1):
    """Fused form_schur_system + pcg: d_G_dense -> G^-1, d_lambda in/out."""
    s, c, N = state_size, control_size, knot_points
    z = sizes(s, c, N)
    f, sfx, cT = _prec(d_G_dense, "d_G_dense")
    rc = getattr(lib(), "mpcgpu_b200_kkt_solve_" + sfx)(
        ctypes.c_uint32(s), ctypes.c_uint32(c), ctypes.c_uint32(N), ctypes.c_uint32(batch),
        _ptr(d_G_dense, f, "d_G_dense", z["G"] * batch),
        _ptr(d_C_dense, f, "d_C_dense", z["C"] * batch) if N > 1 else None,
        _ptr(d_g, f, "d_g", z["g"] * batch), _ptr(d_c, f, "d_c", z["c"] * batch),
        cT(rho), _ptr(d_lambda, f, "d_lambda", z["lambda"] * batch),
        ctypes.c_uint32(config.pcg_max_iter), cT(config.pcg_exit_tol),
        _ptr(d_pcg_iters, torch.int32, "d_pcg_iters", batch),
        _ptr(d_pcg_exit, d_pcg_exit.dtype, "d_pcg_exit", batch), None, _stream())
    _check(rc, "kkt_solve")


def kkt_solve_dz(state_size, control_size, knot_points, d_G_dense, d_C_dense, d_g, d_c, rho,
                 d_lambda, d_dz, d_pcg_iters, d_pcg_exit, config: PcgConfig, batch=1):
    """kkt_solve with compute_dz (include/common/dz.cuh:124-133) as the kernel's epilogue."""
    s, c, N = state_size, control_size, knot_points
    z = sizes(s, c, N)
    f, sfx, cT = _prec(d_G_dense, "d_G_dense")
    rc = getattr(lib(), "mpcgpu_b200_kkt_solve_dz_" + sfx)(
        ctypes.c_uint32(s), ctypes.c_uint32(c), ctypes.c_uint32(N), ctypes.c_uint32(batch),
        _ptr(d_G_dense, f, "d_G_dense", z["G"] * batch),
        _ptr(d_C_dense, f, "d_C_dense", z["C"] * batch) if N > 1 else None,
        _ptr(d_g, f, "d_g", z["g"] * batch), _ptr(d_c, f, "d_c", z["c"] * batch),
        cT(rho), _ptr(d_lambda, f, "d_lambda", z["lambda"] * batch),
        _ptr(d_dz, f, "d_dz", z["g"] * batch),
        ctypes.c_uint32(config.pcg_max_iter), cT(config.pcg_exit_tol),
        _ptr(d_pcg_iters, torch.int32, "d_pcg_iters", batch),
        _ptr(d_pcg_exit, d_pcg_exit.dtype, "d_pcg_exit", batch), None, _stream())
    _check(rc, "kkt_solve_dz")


def compute_dz(state_size, control_size, knot_points, d_G_dense, d_C_dense, d_g_val, d_lambda, d_dz, batch=1):
    """compute_dz of include/common/dz.cuh:124-133 (same argument order): d_G_dense holds G^-1."""
    s, c, N = state_size, control_size, knot_points
    z = sizes(s, c, N)
    f, sfx, cT = _prec(d_G_dense, "d_G_dense")
    rc = getattr(lib(), "mpcgpu_b200_dz_" + sfx)(
        ctypes.c_uint32(s), ctypes.c_uint32(c), ctypes.c_uint32(N), ctypes.c_uint32(batch),
        _ptr(d_G_dense, f, "d_G_dense", z["G"] * batch),
        _ptr(d_C_dense, f, "d_C_dense", z["C"] * batch) if N > 1 else None,
        _ptr(d_g_val, f, "d_g_val", z["g"] * batch), _ptr(d_lambda, f, "d_lambda", z["lambda"] * batch),
        _ptr(d_dz, f, "d_dz", z["g"] * batch), _stream())
    _check(rc, "compute_dz")


def launch_count():
    return int(lib().mpcgpu_b200_launch_count())


class HostSolver:
    """Solver object over host (ideally pinned) buffers: H2D, fused solve and D2H are
    pipelined in chunks inside the library (mpcgpu_b200_solver_* in the C ABI)."""

    def __init__(self, state_size, control_size, knot_points, max_batch, device=0):
        self.s, self.c, self.N, self.max_batch = state_size, control_size, knot_points, max_batch
        self._h = ctypes.c_void_p()
        rc = lib().mpcgpu_b200_solver_create(
            ctypes.byref(self._h), ctypes.c_int(device), ctypes.c_uint32(state_size),
            ctypes.c_uint32(control_size), ctypes.c_uint32(knot_points), ctypes.c_uint32(max_batch))
        _check(rc, "solver_create")

    @staticmethod
    def _hp(t, dtype, n, what):
        if t is None:
            return None
        if t.is_cuda or t.dtype != dtype or not t.is_contiguous() or t.numel() < n:
            raise MpcGpuError("%s must be a contiguous host %s tensor with >= %d elements" % (what, dtype, n))
        return ctypes.c_void_p(t.data_ptr())

    def solve(self, h_G, h_C, h_g, h_c, rho, h_lambda, h_iters, h_exit, config: PcgConfig,
              batch, h_Ginv=None, h_lambda0="inplace"):
        """h_lambda0: warm start; "inplace" (default) reads it from h_lambda, None = cold start
        (lambda0 = 0, nothing copied in), or a separate host tensor."""
        z = sizes(self.s, self.c, self.N)
        f = torch.float32
        lam0 = h_lambda if isinstance(h_lambda0, str) else h_lambda0
        rc = lib().mpcgpu_b200_solver_solve_host_io_f32(
            self._h, ctypes.c_uint32(batch),
            self._hp(h_G, f, z["G"] * batch, "h_G"), self._hp(h_C, f, z["C"] * batch, "h_C"),
            self._hp(h_g, f, z["g"] * batch, "h_g"), self._hp(h_c, f, z["c"] * batch, "h_c"),
            ctypes.c_float(rho), self._hp(lam0, f, z["lambda"] * batch, "h_lambda0"),
            self._hp(h_lambda, f, z["lambda"] * batch, "h_lambda"),
            ctypes.c_uint32(config.pcg_max_iter), ctypes.c_float(config.pcg_exit_tol),
            self._hp(h_iters, torch.int32, batch, "h_iters"),
            self._hp(h_exit, torch.uint8, batch, "h_exit"),
            self._hp(h_Ginv, f, z["G"] * batch, "h_Ginv"))
        _check(rc, "solver_solve_host")

    def close(self):
        if self._h:
            lib().mpcgpu_b200_solver_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---- long horizon split by knot range across GPUs (BASELINE.json configs[3]) ----------------------
def knot_range(knot_points, rank, world):
    """Knots [lo, hi) owned by `rank`: equal contiguous ranges (the kernels need 16, 32, 64, 128,
    256 or 512 knots per rank)."""
    if knot_points % world:
        raise MpcGpuError("knot_points %d not divisible by world %d" % (knot_points, world))
    n = knot_points // world
    if n not in (16, 32, 64, 128, 256, 512):
        raise MpcGpuError("knots per rank must be 16..512 in powers of two, got %d" % n)
    return rank * n, (rank + 1) * n


def schur_slice(knot_points, rank, world):
    """Extended knot range [lo, hi) a rank feeds to form_schur_system so that the S / Pinv /
    gamma block rows of ITS knots come out exactly as in the full-horizon computation, with no
    communication: two extra knots to the left (theta^-1 of knot lo-1 needs phi of lo-1, which
    needs Q of lo-2) and one to the right (S-right / Pinv-right of the last owned knot).
    Returns (lo_ext, hi_ext, offset of the first owned knot inside the slice)."""
    lo, hi = knot_range(knot_points, rank, world)
    lo_e, hi_e = max(lo - 2, 0), min(hi + 1, knot_points)
    return lo_e, hi_e, lo - lo_e


def schur_rows_local(state_size, control_size, knot_points, rank, world, G, C, g, c, rho):
    """S / Pinv / gamma block rows of the knots `rank` owns, formed on THIS device from its slice of the full-horizon
    inputs (host or device 1-D float32 tensors G, C, g, c of the whole horizon) with no communication.  Ranges longer
    than one form_schur_system call can take (512 knots incl. the 2 + 1 halo knots of :func:`schur_slice`) are
    formed in sub-ranges; every row is bit-identical to the full-horizon computation.
    Returns (S_local, Pinv_local, gamma_local) as contiguous device tensors."""
    s, cc, N = state_size, control_size, knot_points
    lo, hi = knot_range(N, rank, world)
    dev = torch.device("cuda", torch.cuda.current_device())
    nG, nC, ng = s * s + cc * cc, s * s + s * cc, s + cc
    S_l = torch.empty(3 * s * s * (hi - lo), device=dev)
    P_l = torch.empty_like(S_l)
    g_l = torch.empty(s * (hi - lo), device=dev)
    step = 500
    for a in range(lo, hi, step):
        b = min(a + step, hi)
        lo_e, hi_e = max(a - 2, 0), min(b + 1, N)
        ne, off = hi_e - lo_e, a - lo_e
        z = sizes(s, cc, ne)
        cut = lambda t, per, n: t[lo_e * per: lo_e * per + n].to(dev, non_blocking=True).contiguous()[None].clone()
        Ge, Ce, ge, ce = cut(G, nG, z["G"]), cut(C, nC, z["C"]), cut(g, ng, z["g"]), cut(c, s, z["c"])
        Se = torch.empty((1, z["S"]), device=dev); Pe = torch.empty_like(Se); gme = torch.empty((1, z["gamma"]), device=dev)
        form_schur_system(s, cc, ne, Ge, Ce, ge, ce, Se, Pe, gme, rho)
        w = 3 * s * s
        S_l[(a - lo) * w:(b - lo) * w] = Se[0, off * w:(off + b - a) * w]
        P_l[(a - lo) * w:(b - lo) * w] = Pe[0, off * w:(off + b - a) * w]
        g_l[(a - lo) * s:(b - lo) * s] = gme[0, off * s:(off + b - a) * s]
    return S_l, P_l, g_l


def mailbox_bytes():
    lib().mpcgpu_b200_knotsplit_mailbox_bytes.restype = ctypes.c_size_t
    return int(lib().mpcgpu_b200_knotsplit_mailbox_bytes())


def knotsplit_pcg(state_size, knots_local, ranks_here, rank_first, world, epoch, mailboxes, d_S, d_Pinv,
                  d_gamma, d_lambda, d_pcg_iters, d_pcg_exit, config: PcgConfig):
    """PCG on a knot-range-split horizon.  `mailboxes`: `world` integers (device addresses of every
    rank's mailbox as seen from THIS device).  d_S / d_Pinv / d_gamma / d_lambda hold the slices of
    the `ranks_here` ranks resident on this device, in rank order (ranks_here > 1: single-GPU
    emulation)."""
    s, n, f = state_size, knots_local * ranks_here, torch.float32
    arr = (ctypes.c_void_p * world)(*[ctypes.c_void_p(int(m)) for m in mailboxes])
    rc = lib().mpcgpu_b200_knotsplit_pcg_f32(
        ctypes.c_uint32(s), ctypes.c_uint32(knots_local), ctypes.c_uint32(ranks_here),
        ctypes.c_uint32(rank_first), ctypes.c_uint32(world), ctypes.c_uint32(epoch), arr,
        _ptr(d_S, f, "d_S", 3 * s * s * n), _ptr(d_Pinv, f, "d_Pinv", 3 * s * s * n),
        _ptr(d_gamma, f, "d_gamma", s * n), _ptr(d_lambda, f, "d_lambda", s * n),
        ctypes.c_uint32(config.pcg_max_iter), ctypes.c_float(config.pcg_exit_tol),
        _ptr(d_pcg_iters, torch.int32, "d_pcg_iters", ranks_here),
        _ptr(d_pcg_exit, d_pcg_exit.dtype, "d_pcg_exit", ranks_here), _stream())
    _check(rc, "knotsplit_pcg")


def epoch_stride(config: PcgConfig):
    """Epoch increment between consecutive solves of a group: odd (alternates the mailbox half) and larger than
    the number of tags a solve can use."""
    lib().mpcgpu_b200_knotsplit_epoch_stride.restype = ctypes.c_uint32
    return int(lib().mpcgpu_b200_knotsplit_epoch_stride(ctypes.c_uint32(int(config.pcg_max_iter))))


def knotsplit_last_timeout():
    """{rank, expected_tag, seen_tag} of a peer poll of this process that timed out, else None."""
    out = (ctypes.c_uint32 * 4)()
    if lib().mpcgpu_b200_knotsplit_last_timeout(out):
        return {"rank": int(out[1]), "expected_tag": int(out[2]), "seen_tag": int(out[3])}
    return None


class KnotSplitGroup:
    """One process per GPU (torch.distributed initialised): allocates this rank's mailbox, maps
    the peers' mailboxes through CUDA IPC and keeps the solve epoch in step on all ranks."""

    def __init__(self, group=None):
        import torch.distributed as dist
        self.dist, self.group = dist, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self._mail = ctypes.c_void_p()
        _check(lib().mpcgpu_b200_knotsplit_mailbox_create(ctypes.byref(self._mail)), "mailbox_create")
        h = (ctypes.c_ubyte * 64)()
        _check(lib().mpcgpu_b200_ipc_export(self._mail, h), "ipc_export")
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(h), group=group)
        self.ptrs, self._opened = [], []
        for r, hb in enumerate(handles):
            if r == self.rank:
                self.ptrs.append(self._mail.value)
            else:
                p = ctypes.c_void_p()
                buf = (ctypes.c_ubyte * 64).from_buffer_copy(hb)
                _check(lib().mpcgpu_b200_ipc_open(buf, ctypes.byref(p)), "ipc_open")
                self.ptrs.append(p.value)
                self._opened.append(p)
        self.epoch = 0
        dist.barrier(group=group)

    def pcg(self, state_size, knots_local, d_S, d_Pinv, d_gamma, d_lambda, d_iters, d_exit, config):
        knotsplit_pcg(state_size, knots_local, 1, self.rank, self.world, self.epoch, self.ptrs, d_S, d_Pinv,
                      d_gamma, d_lambda, d_iters, d_exit, config)
        self.epoch += epoch_stride(config)

    def pingpong_us(self, reps=2000):
        """One-way latency (us) of ONE tagged 8-byte peer store + local poll between ranks 0 and 1 over NVLink:
        the floor of a knot-split exchange.  Collective: every rank of the group calls it; ranks >= 2 idle."""
        out = torch.zeros(1, dtype=torch.int64, device="cuda")
        self.epoch += 2 * reps + 17
        if self.rank < 2 and self.world >= 2:
            peer = self.ptrs[1 - self.rank]
            _check(lib().mpcgpu_b200_knotsplit_pingpong(
                ctypes.c_uint32(self.rank), ctypes.c_void_p(self.ptrs[self.rank]), ctypes.c_void_p(peer),
                ctypes.c_uint32(reps), ctypes.c_uint32(self.epoch), ctypes.c_void_p(out.data_ptr()), _stream()), "knotsplit_pingpong")
        torch.cuda.synchronize()
        self.epoch += 2 * reps + 18          # keeps the odd stride of the solves that follow
        self.dist.barrier(group=self.group)
        t = out.to(torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX, group=self.group)
        return float(t[0]) / (2.0 * reps) * 1e-3

    def close(self):
        torch.cuda.synchronize()
        self.dist.barrier(group=self.group)
        for p in self._opened:
            lib().mpcgpu_b200_ipc_close(p)
        self._opened = []
        self.dist.barrier(group=self.group)
        if self._mail:
            lib().mpcgpu_b200_knotsplit_mailbox_destroy(self._mail)
            self._mail = ctypes.c_void_p()
