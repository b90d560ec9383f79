"""Seeded synthetic KKT blocks (G, C, g, c) with the structure the reference's
``generate_kkt_submatrices`` produces for the Kuka IIWA end-effector tracking problem.

Structure sources (reference checkout):
  * packing of d_G / d_C / d_g / d_c ........ include/common/kkt.cuh:106-117,154-160
  * A = I + dt*[[0,I],[dqdd/dq, dqdd/dqd]], B = dt*[[0],[Minv]]
                                              include/common/integrator.cuh:67-79
  * Q = blkdiag(grad grad^T, QD_COST*I), R = R_COST*I
                                              include/dynamics/iiwa/iiwa_eepos_plant.cuh:329-376
  * QD_COST, R_COST ......................... include/common/settings.cuh:90-100
  * d_C stores Abar = -A and Bbar = -B ...... include/common/kkt.cuh:115-116,158-159
  * rho = 1e-3, timestep 1/64 ............... include/mpcsim.cuh:256, examples/track_iiwa_pcg.cu:19

The real blocks need the GRiD rigid-body dynamics (out of scope); this
generator reproduces their sparsity, scaling and conditioning (recipe: SURVEY 8(d)).
"""
import numpy as np

BASE_SEED = 0x4D504347  # "MPCG"
QD_COST = 1e-4
RHO = 1e-3
DT = 1.0 / 64.0


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
    }


def r_cost(N):
    return 1e-3 if N == 64 else 1e-4  # settings.cuh:90-96


def _base_trajectory(rng, N, s, c):
    nq = s // 2
    v = rng.standard_normal((N, nq)) * 0.3
    L = np.tril(rng.standard_normal((N, nq, nq)))
    Minv = L @ L.transpose(0, 2, 1) + 0.5 * np.eye(nq)
    Jq = rng.standard_normal((N, nq, nq)) * 2.0
    Jqd = rng.standard_normal((N, nq, nq)) * 0.5
    e = rng.standard_normal(N) * 0.05
    qd = rng.standard_normal((N, nq)) * 0.5
    u = rng.standard_normal((N, c)) * 5.0
    cres = rng.standard_normal((N, s)) * 1e-3
    cres[0] = 0.0
    return dict(v=v, Minv=Minv, Jq=Jq, Jqd=Jqd, e=e, qd=qd, u=u, cres=cres)


def gen_kkt(N, batch=1, seed=BASE_SEED, perturb=0.02, s=14, c=7, dtype=np.float32):
    """Returns dict(G, C, g, c, rho) with arrays of shape [batch, size] (float64 masters
    are rounded once to `dtype`, so CPU oracle and GPU see bit-identical inputs).

    Instance 0 is the base trajectory; instance i>0 applies multiplicative noise
    ``1 + N(0, perturb)`` to the dynamics Jacobians, Minv, gradients and residuals
    ("perturbed start states", BASELINE.json configs[4])."""
    assert s % 2 == 0 and c <= s // 2
    nq = s // 2
    rng = np.random.Generator(np.random.PCG64(seed))
    base = _base_trajectory(rng, N, s, c)
    rc = r_cost(N)
    sz = sizes(s, c, N)
    G = np.zeros((batch, sz["G"]))
    C = np.zeros((batch, sz["C"]))
    g = np.zeros((batch, sz["g"]))
    cv = np.zeros((batch, sz["c"]))
    Gset, Cset, gset = s * s + c * c, s * s + s * c, s + c

    def noise(shape, b):
        if b == 0 or perturb == 0.0:
            return np.ones(shape)
        return 1.0 + perturb * prng.standard_normal(shape)

    for b in range(batch):
        prng = np.random.Generator(np.random.PCG64(seed + 1000003 * b + 1))
        v = base["v"] * noise(base["v"].shape, b)
        Minv = base["Minv"] * noise((N, 1, 1), b)          # keeps symmetry / SPD
        Jq = base["Jq"] * noise(base["Jq"].shape, b)
        Jqd = base["Jqd"] * noise(base["Jqd"].shape, b)
        e = base["e"] * noise(base["e"].shape, b)
        qd = base["qd"] * noise(base["qd"].shape, b)
        u = base["u"] * noise(base["u"].shape, b)
        cres = base["cres"] * noise(base["cres"].shape, b)
        for k in range(N):
            Q = np.zeros((s, s))
            Q[:nq, :nq] = np.outer(v[k], v[k])
            Q[nq:, nq:] = QD_COST * np.eye(nq)
            G[b, k * Gset:k * Gset + s * s] = Q.T.reshape(-1)   # column-major
            g[b, k * gset:k * gset + nq] = v[k] * e[k]
            g[b, k * gset + nq:k * gset + s] = QD_COST * qd[k]
            if k < N - 1:
                G[b, k * Gset + s * s:(k + 1) * Gset] = (rc * np.eye(c)).reshape(-1)
                g[b, k * gset + s:(k + 1) * gset] = rc * u[k]
                A = np.eye(s)
                A[:nq, nq:] += DT * np.eye(nq)
                A[nq:, :nq] += DT * Jq[k]
                A[nq:, nq:] += DT * Jqd[k]
                B = np.zeros((s, c))
                B[nq:, :] = DT * Minv[k][:, :c]
                C[b, k * Cset:k * Cset + s * s] = (-A).T.reshape(-1)
                C[b, k * Cset + s * s:(k + 1) * Cset] = (-B).T.reshape(-1)
            cv[b, k * s:(k + 1) * s] = cres[k]
    return {
        "G": G.astype(dtype), "C": C.astype(dtype), "g": g.astype(dtype), "c": cv.astype(dtype),
        "rho": RHO, "s": s, "c_size": c, "N": N, "batch": batch,
    }


def gen_kkt_fast(N, batch, seed=BASE_SEED, perturb=0.02, s=14, c=7, dtype=np.float32, perturb_seed=None):
    """Vectorised generator for large batches (bench): same structure as :func:`gen_kkt`
    but one RNG stream for the whole batch (not bit-identical to gen_kkt).
    perturb_seed: draw the per-instance perturbations from their own stream, so that several shards
    (one per GPU) share the base trajectory `seed` but hold different perturbed instances of it."""
    nq = s // 2
    rng = np.random.Generator(np.random.PCG64(seed))
    base = _base_trajectory(rng, N, s, c)
    if perturb_seed is not None:
        rng = np.random.Generator(np.random.PCG64([int(seed), int(perturb_seed)]))
    rc = r_cost(N)
    sz = sizes(s, c, N)
    Gset, Cset, gset = s * s + c * c, s * s + s * c, s + c

    def pert(x, shape=None):
        shp = (batch,) + (x.shape if shape is None else shape)
        n = 1.0 + perturb * rng.standard_normal(shp)
        n[0] = 1.0
        return x[None] * n

    v = pert(base["v"])
    Minv = pert(base["Minv"], (N, 1, 1))
    Jq, Jqd = pert(base["Jq"]), pert(base["Jqd"])
    e, qd, u, cres = pert(base["e"]), pert(base["qd"]), pert(base["u"]), pert(base["cres"])
    Q = np.zeros((batch, N, s, s))
    Q[:, :, :nq, :nq] = v[..., :, None] * v[..., None, :]
    Q[:, :, nq:, nq:] = QD_COST * np.eye(nq)
    R = np.broadcast_to(rc * np.eye(c), (batch, N, c, c))
    Gfull = np.concatenate([Q.transpose(0, 1, 3, 2).reshape(batch, N, s * s),
                            R.reshape(batch, N, c * c)], axis=2).reshape(batch, N * Gset)
    G = Gfull[:, :sz["G"]]
    gq = np.concatenate([v * e[..., None], QD_COST * qd, rc * u], axis=2).reshape(batch, N * gset)
    g = gq[:, :sz["g"]]
    A = np.zeros((batch, N, s, s))
    A[:] = np.eye(s)
    A[:, :, :nq, nq:] += DT * np.eye(nq)
    A[:, :, nq:, :nq] += DT * Jq
    A[:, :, nq:, nq:] += DT * Jqd
    B = np.zeros((batch, N, s, c))
    B[:, :, nq:, :] = DT * Minv[..., :, :c]
    Cfull = np.concatenate([(-A).transpose(0, 1, 3, 2).reshape(batch, N, s * s),
                            (-B).transpose(0, 1, 3, 2).reshape(batch, N, s * c)], axis=2)
    C = Cfull[:, :N - 1].reshape(batch, (N - 1) * Cset)
    cv = cres.reshape(batch, N * s)
    return {
        "G": np.ascontiguousarray(G, dtype), "C": np.ascontiguousarray(C, dtype),
        "g": np.ascontiguousarray(g, dtype), "c": np.ascontiguousarray(cv, dtype),
        "rho": RHO, "s": s, "c_size": c, "N": N, "batch": batch,
    }
