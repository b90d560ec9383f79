"""TEST INFRASTRUCTURE ONLY -- numpy/ctypes front end of the CPU oracle.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package.  The product
(``mpcgpu_b200``) never does: it fails loudly when its CUDA library is missing.

Functions mirror the reference's hot-path entry points (all arrays column-major small
blocks packed exactly like the reference's device buffers, see
``include/pcg/sqp.cuh:54-58`` for the sizes):

* :func:`schur`          -- ``form_schur_system`` (``include/pcg/linsys_setup.cuh:620-657``)
* :func:`pcg`            -- ``pcg<T,S,N>`` (call site ``include/pcg/sqp.cuh:142-166,275``)
* :func:`qdldl_solve`    -- ``qdldl_solve_schur`` (``include/qdldl/sqp.cuh:22-49``) using the
  REFERENCE's own QDLDL build from ``oracle/_ref``
* :func:`dz`             -- ``compute_dz`` (``include/common/dz.cuh``)
"""
import ctypes
import os

import numpy as np

from . import build as _build

_HERE = os.path.dirname(os.path.abspath(__file__))
_libs = {}


def _lib(name):
    if not _libs:
        paths = _build.build(reference_cuda=False)
        for k, p in paths.items():
            _libs[k] = ctypes.CDLL(p)
    return _libs[name]


def _dt(dtype):
    dtype = np.dtype(dtype)
    if dtype == np.float32:
        return "f32", ctypes.c_float
    if dtype == np.float64:
        return "f64", ctypes.c_double
    raise TypeError(dtype)


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def sizes(s, c, N):
    """Element counts of the reference buffers (include/pcg/sqp.cuh:54-58,111,130)."""
    return {
        "G": (s * s + c * c) * N - c * c,
        "C": (s * s + s * c) * (N - 1),
        "g": (s + c) * N - c,
        "c": s * N,
        "S": 3 * s * s * N,
        "Pinv": 3 * s * s * N,
        "gamma": s * N,
        "lambda": s * N,
        "dz": (s + c) * N - c,
        "nnz": (N - 1) * s * s + N * ((s + 1) * s // 2),
    }


def schur(s, c, N, G, C, g, cvec, rho, dtype=np.float32):
    """Returns (S, Pinv, gamma, Ginv); inputs are not modified."""
    suf, ct = _dt(dtype)
    G = np.array(G, dtype=dtype, copy=True)
    C = np.ascontiguousarray(C, dtype=dtype)
    g = np.ascontiguousarray(g, dtype=dtype)
    cvec = np.ascontiguousarray(cvec, dtype=dtype)
    sz = sizes(s, c, N)
    assert G.size == sz["G"] and C.size == sz["C"] and g.size == sz["g"] and cvec.size == sz["c"]
    S = np.zeros(sz["S"], dtype)
    Pinv = np.zeros(sz["S"], dtype)
    gamma = np.zeros(sz["gamma"], dtype)
    fn = getattr(_lib("oracle"), "oracle_schur_" + suf)
    fn.restype = None
    fn(ctypes.c_int(s), ctypes.c_int(c), ctypes.c_int(N), _p(G), _p(C), _p(g), _p(cvec),
       ct(rho), _p(S), _p(Pinv), _p(gamma))
    return S, Pinv, gamma, G


def bdmv(s, N, M, x):
    dtype = M.dtype
    suf, _ = _dt(dtype)
    x = np.ascontiguousarray(x, dtype=dtype)
    y = np.zeros(s * N, dtype)
    fn = getattr(_lib("oracle"), "oracle_bdmv_" + suf)
    fn.restype = None
    fn(ctypes.c_int(s), ctypes.c_int(N), _p(np.ascontiguousarray(M)), _p(x), _p(y))
    return y


def pcg(s, N, S, Pinv, gamma, lambda0, max_iter, exit_tol, dtype=np.float32, mode=0):
    """Returns (lambda, iters, max_iter_exit).  mode=1: alternate (equally legitimate)
    floating-point evaluation order, used to measure the fp32 parity noise floor."""
    suf, ct = _dt(dtype)
    S = np.ascontiguousarray(S, dtype=dtype)
    Pinv = np.ascontiguousarray(Pinv, dtype=dtype)
    gamma = np.ascontiguousarray(gamma, dtype=dtype)
    lam = np.array(lambda0, dtype=dtype, copy=True)
    iters = ctypes.c_uint(0)
    flag = ctypes.c_ubyte(0)
    fn = getattr(_lib("oracle"), "oracle_pcg_mode_" + suf)
    fn.restype = None
    fn(ctypes.c_int(s), ctypes.c_int(N), _p(S), _p(Pinv), _p(gamma), _p(lam),
       ctypes.c_uint(max_iter), ct(exit_tol), ctypes.byref(iters), ctypes.byref(flag),
       ctypes.c_int(mode))
    return lam, int(iters.value), bool(flag.value)


def pcg_cg1(s, N, S, Pinv, gamma, lambda0, max_iter, exit_tol, dtype=np.float32):
    """Single-reduction PCG (oracle_pcg_cg1): the checker of the product's PCG variant 1.
    Returns (lambda, iters, max_iter_exit)."""
    suf, ct = _dt(dtype)
    S = np.ascontiguousarray(S, dtype=dtype)
    Pinv = np.ascontiguousarray(Pinv, dtype=dtype)
    gamma = np.ascontiguousarray(gamma, dtype=dtype)
    lam = np.array(lambda0, dtype=dtype, copy=True)
    iters = ctypes.c_uint(0)
    flag = ctypes.c_ubyte(0)
    fn = getattr(_lib("oracle"), "oracle_pcg_cg1_" + suf)
    fn.restype = None
    fn(ctypes.c_int(s), ctypes.c_int(N), _p(S), _p(Pinv), _p(gamma), _p(lam),
       ctypes.c_uint(max_iter), ct(exit_tol), ctypes.byref(iters), ctypes.byref(flag))
    return lam, int(iters.value), bool(flag.value)


def pcg_sprec(s, N, S, Pinv, gamma, lambda0, max_iter, exit_tol, dtype=np.float32):
    """PCG with S p carried by a recurrence (oracle_pcg_sprec): the checker of the product's PCG variant 2.
    Returns (lambda, iters, max_iter_exit)."""
    suf, ct = _dt(dtype)
    S = np.ascontiguousarray(S, dtype=dtype)
    Pinv = np.ascontiguousarray(Pinv, dtype=dtype)
    gamma = np.ascontiguousarray(gamma, dtype=dtype)
    lam = np.array(lambda0, dtype=dtype, copy=True)
    iters = ctypes.c_uint(0)
    flag = ctypes.c_ubyte(0)
    fn = getattr(_lib("oracle"), "oracle_pcg_sprec_" + suf)
    fn.restype = None
    fn(ctypes.c_int(s), ctypes.c_int(N), _p(S), _p(Pinv), _p(gamma), _p(lam),
       ctypes.c_uint(max_iter), ct(exit_tol), ctypes.byref(iters), ctypes.byref(flag))
    return lam, int(iters.value), bool(flag.value)


def pcg_noise_floor(s, N, S32, P32, g32, lambda0, iters):
    """How far correct fp32 implementations of the SAME `iters` PCG iterations drift apart on
    this input: max of (plain-order fp32 vs alternate-order fp32) and (fp32 vs fp64), relative
    to ||lambda||.  The fp64 run uses the identical fp32 inputs and iteration count."""
    a, _, _ = pcg(s, N, S32, P32, g32, lambda0, iters, 0.0, np.float32, mode=0)
    b, _, _ = pcg(s, N, S32, P32, g32, lambda0, iters, 0.0, np.float32, mode=1)
    c, _, _ = pcg(s, N, np.asarray(S32, np.float64), np.asarray(P32, np.float64),
                  np.asarray(g32, np.float64), lambda0, iters, 0.0, np.float64, mode=0)
    nrm = max(np.linalg.norm(c), 1e-300)
    return max(np.linalg.norm(a - b), np.linalg.norm(a - c), np.linalg.norm(b - c)) / nrm


def csc_pattern(s, N):
    sz = sizes(s, c=0, N=N)
    col_ptr = np.zeros(s * N + 1, np.int32)
    row_ind = np.zeros(sz["nnz"], np.int32)
    fn = _lib("oracle").oracle_csc_pattern
    fn.restype = None
    fn(ctypes.c_int(s), ctypes.c_int(N), _p(col_ptr), _p(row_ind))
    return col_ptr, row_ind


def csc_values(s, N, S):
    dtype = S.dtype
    suf, _ = _dt(dtype)
    val = np.zeros(sizes(s, 0, N)["nnz"], dtype)
    fn = getattr(_lib("oracle"), "oracle_csc_values_" + suf)
    fn.restype = None
    fn(ctypes.c_int(s), ctypes.c_int(N), _p(np.ascontiguousarray(S)), _p(val))
    return val


def qdldl_factor_solve(Ap, Ai, Ax, b):
    """etree + factor + solve with the REFERENCE's QDLDL (oracle/_ref), dtype of Ax picks
    the float (reference configuration) or double build.  Mirrors
    qdldl/tests/qdldl_tester.c:63-159 and include/qdldl/sqp.cuh:22-49,193.
    Returns (x, status) with status = QDLDL_etree/QDLDL_factor return (<0 is failure)."""
    dtype = Ax.dtype
    suf, _ = _dt(dtype)
    lib = _lib("qdldl_" + suf)
    n = len(Ap) - 1
    Ap = np.ascontiguousarray(Ap, np.int32)
    Ai = np.ascontiguousarray(Ai, np.int32)
    Ax = np.ascontiguousarray(Ax)
    iwork = np.zeros(3 * n, np.int32)
    Lnz = np.zeros(n, np.int32)
    etree = np.zeros(n, np.int32)
    lib.QDLDL_etree.restype = ctypes.c_int
    sumLnz = lib.QDLDL_etree(ctypes.c_int(n), _p(Ap), _p(Ai), _p(iwork), _p(Lnz), _p(etree))
    if sumLnz < 0:
        return None, int(sumLnz)
    Lp = np.zeros(n + 1, np.int32)
    Li = np.zeros(max(sumLnz, 1), np.int32)
    Lx = np.zeros(max(sumLnz, 1), dtype)
    D = np.zeros(n, dtype)
    Dinv = np.zeros(n, dtype)
    bwork = np.zeros(n, np.uint8)
    fwork = np.zeros(n, dtype)
    lib.QDLDL_factor.restype = ctypes.c_int
    st = lib.QDLDL_factor(ctypes.c_int(n), _p(Ap), _p(Ai), _p(Ax), _p(Lp), _p(Li), _p(Lx), _p(D),
                          _p(Dinv), _p(Lnz), _p(etree), _p(bwork), _p(iwork), _p(fwork))
    if st < 0:
        return None, int(st)
    x = np.array(b, dtype=dtype, copy=True)
    lib.QDLDL_solve.restype = None
    lib.QDLDL_solve(ctypes.c_int(n), _p(Lp), _p(Li), _p(Lx), _p(Dinv), _p(x))
    return x, int(st)


def qdldl_solve(s, N, S, gamma):
    """lambda = S^-1 gamma through CSC packing + the reference QDLDL (dtype of S)."""
    S = np.ascontiguousarray(S)
    Ap, Ai = csc_pattern(s, N)
    Ax = csc_values(s, N, S)
    x, st = qdldl_factor_solve(Ap, Ai, Ax, np.asarray(gamma, S.dtype))
    if x is None:
        raise RuntimeError("QDLDL failed: %d" % st)
    return x


def dz(s, c, N, Ginv, C, g, lam, dtype=np.float32):
    suf, _ = _dt(dtype)
    out = np.zeros(sizes(s, c, N)["dz"], dtype)
    fn = getattr(_lib("oracle"), "oracle_dz_" + suf)
    fn.restype = None
    fn(ctypes.c_int(s), ctypes.c_int(c), ctypes.c_int(N),
       _p(np.ascontiguousarray(Ginv, dtype)), _p(np.ascontiguousarray(C, dtype)),
       _p(np.ascontiguousarray(g, dtype)), _p(np.ascontiguousarray(lam, dtype)), _p(out))
    return out


def cpu_kkt_solve_batch(s, c, N, G, C, g, cvec, rho, nthreads, with_schur=True, S=None, gamma=None):
    """The reference's CPU KKT solve (Schur restatement + reference QDLDL, fp32) over a batch
    on `nthreads` host threads.  Returns (lambda[batch, s*N], wall_s, sum_schur_s, sum_qdldl_s)."""
    lib = _lib("cpu_baseline")
    f32 = np.float32
    if with_schur:
        G = np.ascontiguousarray(G, f32)
        batch = G.size // sizes(s, c, N)["G"]
        args = [_p(G), _p(np.ascontiguousarray(C, f32)), _p(np.ascontiguousarray(g, f32)),
                _p(np.ascontiguousarray(cvec, f32))]
        keep = args
        sin = gin = None
    else:
        S = np.ascontiguousarray(S, f32)
        gamma = np.ascontiguousarray(gamma, f32)
        batch = S.size // sizes(s, c, N)["S"]
        args = [None, None, None, None]
        sin, gin = _p(S), _p(gamma)
    lam = np.zeros((batch, s * N), f32)
    a = ctypes.c_double(0)
    b = ctypes.c_double(0)
    fn = lib.cpu_kkt_solve_batch
    fn.restype = ctypes.c_double
    wall = fn(ctypes.c_int(s), ctypes.c_int(c), ctypes.c_int(N), ctypes.c_int(batch),
              ctypes.c_int(nthreads), ctypes.c_int(1 if with_schur else 0), *args,
              ctypes.c_float(rho), sin, gin, _p(lam), ctypes.byref(a), ctypes.byref(b))
    if wall < 0:
        raise RuntimeError("cpu_kkt_solve_batch failed: %r" % wall)
    return lam, float(wall), float(a.value), float(b.value)


class CpuPool:
    """Persistent host thread pool around the REFERENCE's QDLDL (oracle/_ref/libcpu_baseline.so): threads,
    QDLDL workspaces, CSC pattern and elimination tree live as long as the object (the reference computes
    the tree once, include/qdldl/sqp.cuh:193)."""

    def __init__(self, s, c, N, nthreads):
        self.lib = _lib("cpu_baseline")
        self.s, self.c, self.N, self.nthreads = s, c, N, int(nthreads)
        self.lib.cpu_pool_create.restype = ctypes.c_void_p
        self.lib.cpu_pool_qdldl.restype = ctypes.c_double
        self.lib.cpu_pool_kkt.restype = ctypes.c_double
        self.lib.cpu_pool_destroy.restype = None
        self.h = ctypes.c_void_p(self.lib.cpu_pool_create(ctypes.c_int(s), ctypes.c_int(c), ctypes.c_int(N),
                                                          ctypes.c_int(self.nthreads)))
        if not self.h:
            raise RuntimeError("cpu_pool_create failed")
        self.nnz = sizes(s, c, N)["nnz"]

    def qdldl(self, val, gamma):
        """qdldl_solve_schur (include/qdldl/sqp.cuh:22-49) of every system given as CSC values val[batch, nnz]
        and gamma[batch, s*N].  Returns (lambda, wall_s)."""
        val = np.ascontiguousarray(val, np.float32)
        gamma = np.ascontiguousarray(gamma, np.float32)
        batch = val.size // self.nnz
        lam = np.zeros((batch, self.s * self.N), np.float32)
        wall = self.lib.cpu_pool_qdldl(self.h, ctypes.c_int(batch), _p(val), _p(gamma), _p(lam))
        if wall < 0:
            raise RuntimeError("cpu_pool_qdldl failed: %r" % wall)
        return lam, float(wall)

    def kkt(self, G, C, g, cvec, rho):
        """Schur restatement without preconditioner (include/qdldl/linsys_setup.cuh:12-336) + CSC packing +
        reference QDLDL.  Returns (lambda, wall_s, sum_schur_s, sum_qdldl_s)."""
        f32 = np.float32
        G = np.ascontiguousarray(G, f32)
        batch = G.size // sizes(self.s, self.c, self.N)["G"]
        lam = np.zeros((batch, self.s * self.N), f32)
        a, b = ctypes.c_double(0), ctypes.c_double(0)
        wall = self.lib.cpu_pool_kkt(self.h, ctypes.c_int(batch), _p(G), _p(np.ascontiguousarray(C, f32)),
                                     _p(np.ascontiguousarray(g, f32)), _p(np.ascontiguousarray(cvec, f32)),
                                     ctypes.c_float(rho), _p(lam), ctypes.byref(a), ctypes.byref(b))
        if wall < 0:
            raise RuntimeError("cpu_pool_kkt failed: %r" % wall)
        return lam, float(wall), float(a.value), float(b.value)

    def schur_csc(self, G, C, g, cvec, rho):
        """Preparation of :meth:`qdldl`'s inputs (what the reference's GPU Schur kernel hands to the host):
        returns (val[batch, nnz], gamma[batch, s*N])."""
        f32 = np.float32
        G = np.ascontiguousarray(G, f32)
        batch = G.size // sizes(self.s, self.c, self.N)["G"]
        val = np.zeros((batch, self.nnz), f32)
        gam = np.zeros((batch, self.s * self.N), f32)
        self.lib.cpu_pool_schur_csc.restype = ctypes.c_double
        wall = self.lib.cpu_pool_schur_csc(self.h, ctypes.c_int(batch), _p(G), _p(np.ascontiguousarray(C, f32)),
                                           _p(np.ascontiguousarray(g, f32)), _p(np.ascontiguousarray(cvec, f32)),
                                           ctypes.c_float(rho), _p(val), _p(gam))
        if wall < 0:
            raise RuntimeError("cpu_pool_schur_csc failed: %r" % wall)
        return val, gam

    def close(self):
        if self.h:
            self.lib.cpu_pool_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def qdldl_single_core_us(s, N, S, gamma, reps=200, warmup=20):
    """Median single-core microseconds of the reference's host solve of ONE system (QDLDL_factor + copy +
    QDLDL_solve on precomputed CSC values, float build; include/qdldl/sqp.cuh:22-49)."""
    lib = _lib("cpu_baseline")
    val = csc_values(s, N, np.ascontiguousarray(S, np.float32))
    gamma = np.ascontiguousarray(gamma, np.float32)
    lam = np.zeros(s * N, np.float32)
    lib.cpu_qdldl_single_us.restype = ctypes.c_double
    us = lib.cpu_qdldl_single_us(ctypes.c_int(s), ctypes.c_int(N), _p(val), _p(gamma), ctypes.c_int(reps),
                                 ctypes.c_int(warmup), _p(lam))
    if us < 0:
        raise RuntimeError("cpu_qdldl_single_us failed: %r" % us)
    return float(us), lam
