"""Shared helpers for the parity tests (CPU side)."""
import numpy as np

S, C = 14, 7


def dense_from_bd(s, N, M):
    """Dense (s*N)x(s*N) matrix from the [N][3][s*s] column-major block rows."""
    M = np.asarray(M, np.float64).reshape(N, 3, s, s).transpose(0, 1, 3, 2)  # -> [k][b][row][col]
    D = np.zeros((s * N, s * N))
    for k in range(N):
        if k > 0:
            D[k * s:(k + 1) * s, (k - 1) * s:k * s] = M[k, 0]
        D[k * s:(k + 1) * s, k * s:(k + 1) * s] = M[k, 1]
        if k < N - 1:
            D[k * s:(k + 1) * s, (k + 1) * s:(k + 2) * s] = M[k, 2]
    return D


def dense_schur_definition(s, c, N, G, C, g, cv, rho):
    """Independent fp64 statement of the math in SURVEY 3.2:
       S_stored = -(C Gt^-1 C^T), gamma_stored = -(C Gt^-1 g - c~), Gt = G + rho I, where the
       constraint Jacobian has I on the block diagonal and [Abar_k Bbar_k] on the sub-diagonal.
       Row 0 of the constraint has residual 0 in gamma (d_c[0:s] is never read)."""
    nz = (s + c) * N - c
    Gt = np.zeros((nz, nz))
    gv = np.asarray(g, np.float64)
    Cm = np.zeros((s * N, nz))
    G = np.asarray(G, np.float64)
    Cb = np.asarray(C, np.float64)
    for k in range(N):
        o = k * (s + c)
        Q = G[k * (s * s + c * c):k * (s * s + c * c) + s * s].reshape(s, s).T
        Gt[o:o + s, o:o + s] = Q + rho * np.eye(s)
        if k < N - 1:
            R = G[k * (s * s + c * c) + s * s:(k + 1) * (s * s + c * c)].reshape(c, c).T
            Gt[o + s:o + s + c, o + s:o + s + c] = R + rho * np.eye(c)
        Cm[k * s:(k + 1) * s, o:o + s] = np.eye(s)
        if k > 0:
            op = (k - 1) * (s + c)
            A = Cb[(k - 1) * (s * s + s * c):(k - 1) * (s * s + s * c) + s * s].reshape(s, s).T
            B = Cb[(k - 1) * (s * s + s * c) + s * s:k * (s * s + s * c)].reshape(c, s).T
            Cm[k * s:(k + 1) * s, op:op + s] = A
            Cm[k * s:(k + 1) * s, op + s:op + s + c] = B
    Gi = np.linalg.inv(Gt)
    cres = np.asarray(cv, np.float64).copy()
    cres[:s] = 0.0
    Sd = -(Cm @ Gi @ Cm.T)
    gam = -(Cm @ Gi @ gv - cres)
    return Sd, gam, Gi


def rel(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)


def block_rel(s, N, A, B):
    """max over written blocks of ||A_blk - B_blk||_F / ||B_blk||_F (corner blocks skipped)."""
    A = np.asarray(A, np.float64).reshape(N, 3, s * s)
    B = np.asarray(B, np.float64).reshape(N, 3, s * s)
    worst = 0.0
    for k in range(N):
        for b in range(3):
            if (k == 0 and b == 0) or (k == N - 1 and b == 2):
                continue
            worst = max(worst, np.linalg.norm(A[k, b] - B[k, b]) / max(np.linalg.norm(B[k, b]), 1e-300))
    return worst
