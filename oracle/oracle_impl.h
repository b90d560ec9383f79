/*
 * oracle_impl.h -- TEST INFRASTRUCTURE ONLY (CPU oracle). Not part of the product path.
 *
 * Type-generic body of the CPU restatement of MPCGPU's linear-system hot path.
 * Included twice by mpc_oracle.c with REAL=float / REAL=double and a matching
 * name suffix (F(name) -> name_f32 / name_f64).
 *
 * Every function cites the reference lines it restates (paths relative to the
 * reference checkout).  All matrices are small dense COLUMN-MAJOR blocks.
 *
 * Parity status: the Schur / preconditioner formation and the CSC packing are
 * restated from in-tree reference sources and solved with the REFERENCE's own
 * QDLDL (oracle/_ref, built from qdldl/src/qdldl.c, pinned by QDLDL's 9 KATs).
 * The PCG iteration (oracle_pcg) restates the ABSENT GBD-PCG submodule
 * (A2R-Lab/GBD-PCG @ f0263b84) from its call site include/pcg/sqp.cuh:142-166,275
 * and the published algorithm: for that function "parity unpinned".
 */

/* ---------------------------------------------------------------------------
 * Gauss-Jordan inverse WITHOUT pivoting on the augmented [A | I] matrix with the
 * sliding (DIM+1)-column window.
 *   variant 0: single-matrix form, include/utils/matrix.cuh:120-148
 *              (pvInv = 1/pivot; pivot row *= pvInv; other -= col*pvInv*row)
 *   variant 1: 2-/3-matrix form,   include/utils/matrix.cuh:151-238
 *              (pivot row /= pivot;  other -= col/pivot*row)
 * A: n*n col-major (destroyed, ends as I), Ainv: n*n output.
 * ------------------------------------------------------------------------- */
static void F(gj_invert)(int n, const REAL *A_in, REAL *Ainv, int variant)
{
    REAL M[2 * ORACLE_MAX_DIM * ORACLE_MAX_DIM];
    REAL colv[ORACLE_MAX_DIM], rowv[ORACLE_MAX_DIM + 1];
    int i, j, p, cc;
    for (j = 0; j < n; j++)
        for (i = 0; i < n; i++) {
            M[i + j * n] = A_in[i + j * n];
            M[i + (n + j) * n] = (REAL)(i == j);
        }
    for (p = 0; p < n; p++) {
        REAL piv = M[p + p * n];
        REAL pvInv = (REAL)1 / piv;
        for (i = 0; i < n; i++) colv[i] = M[i + p * n];
        for (cc = 0; cc <= n; cc++) rowv[cc] = M[p + (p + cc) * n];
        for (cc = 0; cc <= n; cc++) {
            for (i = 0; i < n; i++) {
                REAL *e = &M[i + (p + cc) * n];
                if (variant == 0) {
                    if (i == p) *e = *e * pvInv;
                    else        *e = *e - colv[i] * pvInv * rowv[cc];
                } else {
                    if (i == p) *e = *e / piv;
                    else        *e = *e - colv[i] / piv * rowv[cc];
                }
            }
        }
    }
    for (j = 0; j < n * n; j++) Ainv[j] = M[n * n + j];
}

/* out(m x k) = A(m x n) * B(n x k), or A * B^T with B (k x n) when transB.
 * Semantics of glass::gemm<T,TRANSPOSE_B>(m,n,k,alpha,A,B,C) as pinned by the call
 * sites include/pcg/linsys_setup.cuh:398,406,447-455,473-481 (GLASS source absent). */
static void F(gemm)(int m, int n, int k, const REAL *A, const REAL *B, REAL *C, int transB)
{
    int i, j, l;
    for (j = 0; j < k; j++)
        for (i = 0; i < m; i++) {
            REAL acc = (REAL)0;
            for (l = 0; l < n; l++) {
                REAL b = transB ? B[j + l * k] : B[l + j * n];
                acc += A[i + l * m] * b;
            }
            C[i + j * m] = acc;
        }
}

/* include/utils/matrix.cuh:44-55 */
static void F(matvec)(int rows, int cols, const REAL *M, const REAL *v, REAL *out)
{
    int r, c;
    for (r = 0; r < rows; r++) {
        REAL acc = (REAL)0;
        for (c = 0; c < cols; c++) acc += M[r + c * rows] * v[c];
        out[r] = acc;
    }
}

/* block (block_row, block_col) of the [N][3][s*s] block-tridiagonal buffer;
 * store_block_bd / load_block_bd, signatures at
 * include/pcg/linsys_setup_no_precond.cuh:29-49 */
static REAL *F(bd)(REAL *buf, int s, int block_row, int block_col)
{
    return buf + ((size_t)block_row * 3 + block_col) * s * s;
}

/* ---------------------------------------------------------------------------
 * form_schur_system: include/pcg/linsys_setup.cuh:139-562 (phase 1, per block row)
 * and :9-137 (phase 2, SS preconditioner off-diagonals), launched by :565-657.
 * Un-raced semantics: every input is read from the ORIGINAL G.
 *   G  : (s*s+c*c)*N - c*c   in: [Q_k | R_k] per knot; out: [Qt_k^-1 | Rt_k^-1]
 *   C  : (s*s+s*c)*(N-1)     [Abar_k | Bbar_k]  (Abar = -A, Bbar = -B)
 *   g  : (s+c)*N - c         [q_k | r_k]
 *   cv : s*N                 constraint residual (cv[0:s] never read)
 *   S, Pinv : 3*s*s*N ; gamma : s*N
 * Corner blocks S[0][0], S[N-1][2], Pinv[0][0], Pinv[N-1][2] are never written by
 * the reference; the oracle zero-fills them.
 * Pinv == NULL: the Schur formation of the reference's QDLDL slot (form_schur_qdl_kernel,
 * include/qdldl/linsys_setup.cuh:12-336): S, gamma and G^-1 only -- no theta inverse, no
 * phase 2 (that slot never builds a preconditioner).
 * ------------------------------------------------------------------------- */
void F(oracle_schur)(int s, int c, int N, REAL *G, const REAL *C, const REAL *g,
                     const REAL *cv, REAL rho, REAL *S, REAL *Pinv, REAL *gamma)
{
    const int ss = s * s, cc = c * c, sc = s * c;
    const int Gset = ss + cc, Cset = ss + sc, gset = s + c;
    const size_t Gsize = (size_t)Gset * N - cc;
    REAL *G0 = (REAL *)malloc(sizeof(REAL) * Gsize);
    REAL Qk[ORACLE_MAX_DIM * ORACLE_MAX_DIM], Qk_i[ORACLE_MAX_DIM * ORACLE_MAX_DIM];
    REAL Qp[ORACLE_MAX_DIM * ORACLE_MAX_DIM], Qp_i[ORACLE_MAX_DIM * ORACLE_MAX_DIM];
    REAL Rk[ORACLE_MAX_DIM * ORACLE_MAX_DIM], Rk_i[ORACLE_MAX_DIM * ORACLE_MAX_DIM];
    REAL phi[ORACLE_MAX_DIM * ORACLE_MAX_DIM], theta[ORACLE_MAX_DIM * ORACLE_MAX_DIM];
    REAL thetaInv[ORACLE_MAX_DIM * ORACLE_MAX_DIM], BR[ORACLE_MAX_DIM * ORACLE_MAX_DIM];
    REAL tmp[ORACLE_MAX_DIM * ORACLE_MAX_DIM], gam[ORACLE_MAX_DIM];
    REAL t1[ORACLE_MAX_DIM], t2[ORACLE_MAX_DIM];
    int k, i;
    memcpy(G0, G, sizeof(REAL) * Gsize);
    memset(S, 0, sizeof(REAL) * 3 * ss * N);
    if (Pinv) memset(Pinv, 0, sizeof(REAL) * 3 * ss * N);

    /* ---- phase 1, block row 0: linsys_setup.cuh:152-278 ---- */
    memcpy(Qk, G0, sizeof(REAL) * ss);
    for (i = 0; i < s; i++) Qk[i + i * s] += rho;                 /* :180 */
    if (Pinv) for (i = 0; i < ss; i++) F(bd)(Pinv, s, 0, 1)[i] = -Qk[i];    /* :202-210 (before inversion) */
    F(gj_invert)(s, Qk, Qk_i, 1);                                 /* :215-217 (2-matrix form) */
    F(matvec)(s, s, Qk_i, g, gam);                                /* :260-264 */
    for (i = 0; i < ss; i++) F(bd)(S, s, 0, 1)[i] = -Qk_i[i];     /* :249-255 */
    for (i = 0; i < s; i++) gamma[i] = -gam[i];                   /* :273-275 */

    /* ---- phase 1, block rows k >= 1: linsys_setup.cuh:279-560 ---- */
    for (k = 1; k < N; k++) {
        const REAL *Ak = C + (size_t)(k - 1) * Cset;              /* :318 */
        const REAL *Bk = Ak + ss;                                 /* :319 */
        const REAL *qk = g + (size_t)(k - 1) * gset;              /* :323 */
        const REAL *qp = g + (size_t)k * gset;                    /* :324 */
        const REAL *rk = qk + s;                                  /* :325 */
        memcpy(Qk, G0 + (size_t)(k - 1) * Gset, sizeof(REAL) * ss);       /* :320 */
        memcpy(Qp, G0 + (size_t)k * Gset, sizeof(REAL) * ss);             /* :321 */
        memcpy(Rk, G0 + (size_t)(k - 1) * Gset + ss, sizeof(REAL) * cc);  /* :322 */
        for (i = 0; i < s; i++) { Qk[i + i * s] += rho; Qp[i + i * s] += rho; }  /* :329-330 */
        for (i = 0; i < c; i++) Rk[i + i * c] += rho;                             /* :331 */
        F(gj_invert)(s, Qk, Qk_i, 1);                             /* :357-368 (3-matrix form) */
        F(gj_invert)(s, Qp, Qp_i, 1);
        F(gj_invert)(c, Rk, Rk_i, 1);
        memcpy(G + (size_t)(k - 1) * Gset, Qk_i, sizeof(REAL) * ss);          /* :372 */
        memcpy(G + (size_t)(k - 1) * Gset + ss, Rk_i, sizeof(REAL) * cc);     /* :375 */
        if (k == N - 1) memcpy(G + (size_t)k * Gset, Qp_i, sizeof(REAL) * ss); /* :377-380 */

        F(gemm)(s, s, s, Ak, Qk_i, phi, 0);                       /* :398  phi = A Qk^-1 */
        F(gemm)(s, c, c, Bk, Rk_i, BR, 0);                        /* :406  BR  = B R^-1  */
        F(matvec)(s, s, Qp_i, qp, gam);                           /* :411-415 */
        for (i = 0; i < s; i++) gam[i] -= cv[(size_t)k * s + i];  /* :416-418 */
        F(matvec)(s, s, phi, qk, t1);                             /* :422-426 */
        F(matvec)(s, c, BR, rk, t2);                              /* :432-436 */
        for (i = 0; i < s; i++) gam[i] += t2[i] + t1[i];          /* :441-443 */
        F(gemm)(s, s, s, phi, Ak, theta, 1);                      /* :447-455 phi A^T */
        for (i = 0; i < ss; i++) theta[i] += Qp_i[i];             /* :466-468 */
        F(gemm)(s, c, s, BR, Bk, tmp, 1);                         /* :473-481 BR B^T */
        for (i = 0; i < ss; i++) theta[i] += tmp[i];              /* :485-487 */
        for (i = 0; i < ss; i++) F(bd)(S, s, k, 0)[i] = -phi[i];  /* :491-497 */
        for (i = 0; i < ss; i++) F(bd)(S, s, k, 1)[i] = -theta[i];/* :501-507 */
        if (Pinv) {
            F(gj_invert)(s, theta, thetaInv, 0);                  /* :511-513 (single form) */
            for (i = 0; i < ss; i++) F(bd)(Pinv, s, k, 1)[i] = -thetaInv[i]; /* :518-524 */
        }
        for (i = 0; i < s; i++) gamma[(size_t)k * s + i] = -gam[i];      /* :529-532 */
        {   /* :537-557  S[k-1][2] = -(phi)^T */
            int r, q;
            REAL *dst = F(bd)(S, s, k - 1, 2);
            for (q = 0; q < s; q++)
                for (r = 0; r < s; r++) dst[r + q * s] = -phi[q + r * s];
        }
    }

    /* ---- phase 2: linsys_setup.cuh:9-137 ---- */
    for (k = 0; Pinv && k < N; k++) {
        const REAL *thInv_k = F(bd)(Pinv, s, k, 1);
        if (k != 0) {                                             /* :97-115 */
            F(gemm)(s, s, s, thInv_k, F(bd)(S, s, k, 0), tmp, 0);
            F(gemm)(s, s, s, tmp, F(bd)(Pinv, s, k - 1, 1), phi, 0);
            for (i = 0; i < ss; i++) F(bd)(Pinv, s, k, 0)[i] = -phi[i];
        }
        if (k != N - 1) {                                         /* :118-136 */
            int r, q;
            const REAL *pl = F(bd)(S, s, k + 1, 0);               /* loaded transposed :36-43 */
            for (q = 0; q < s; q++)
                for (r = 0; r < s; r++) theta[r + q * s] = pl[q + r * s];
            F(gemm)(s, s, s, thInv_k, theta, tmp, 0);
            F(gemm)(s, s, s, tmp, F(bd)(Pinv, s, k + 1, 1), phi, 0);
            for (i = 0; i < ss; i++) F(bd)(Pinv, s, k, 2)[i] = -phi[i];
        }
    }
    free(G0);
}

/* ---------------------------------------------------------------------------
 * y = M x for the [N][3][s*s] block-tridiagonal layout (GBD-PCG bdmv semantics,
 * source absent; zero-padded at both ends).  The corner blocks are
 * never multiplied.
 * ------------------------------------------------------------------------- */
void F(oracle_bdmv)(int s, int N, const REAL *M, const REAL *x, REAL *y)
{
    int k, i, j;
    const int ss = s * s;
    for (k = 0; k < N; k++) {
        const REAL *L = M + ((size_t)k * 3 + 0) * ss;
        const REAL *D = M + ((size_t)k * 3 + 1) * ss;
        const REAL *R = M + ((size_t)k * 3 + 2) * ss;
        for (i = 0; i < s; i++) {
            REAL acc = (REAL)0;
            if (k > 0)     for (j = 0; j < s; j++) acc += L[i + j * s] * x[(size_t)(k - 1) * s + j];
            for (j = 0; j < s; j++) acc += D[i + j * s] * x[(size_t)k * s + j];
            if (k < N - 1) for (j = 0; j < s; j++) acc += R[i + j * s] * x[(size_t)(k + 1) * s + j];
            y[(size_t)k * s + i] = acc;
        }
    }
}

static REAL F(dot)(size_t n, const REAL *a, const REAL *b)
{
    /* per-knot partial sums then a sum over knots would be the GPU order; the oracle
     * uses a plain left-to-right sum (reduction ORDER is not part of the contract). */
    REAL acc = (REAL)0;
    size_t i;
    for (i = 0; i < n; i++) acc += a[i] * b[i];
    return acc;
}

/* ---------------------------------------------------------------------------
 * Preconditioned conjugate gradient on the stored (negated) Schur system
 *     S lambda = gamma,  preconditioner Pinv (block-tridiagonal, SS).
 * Restates the ABSENT A2R-Lab/GBD-PCG gpu_pcg.cuh @ f0263b84 ("parity unpinned");
 * argument list / outputs pinned by include/pcg/sqp.cuh:142-166,275-277:
 *   lambda is the warm start and the result (include/mpcsim.cuh:227,383),
 *   iters  = number of completed iterations,
 *   max_iter_exit = 1 when the loop ran out of iterations,
 *   exit test: |r . Pinv r| < exit_tol  (ABSOLUTE, on the stored-sign system).
 * ------------------------------------------------------------------------- */
/* Alternate (equally legitimate) fp evaluation order: block products accumulated
 * diag -> left -> right with two interleaved partial sums, dot products as per-knot
 * partial sums reduced by a pairwise tree (the shape of a GPU reduction).  Used by the
 * tests to MEASURE how far two correct fp32 implementations of the same iteration drift
 * apart on a given input (the parity noise floor). */
static void F(bdmv_alt)(int s, int N, const REAL *M, const REAL *x, REAL *y)
{
    int k, i, j;
    const int ss = s * s;
    for (k = 0; k < N; k++) {
        const REAL *L = M + ((size_t)k * 3 + 0) * ss;
        const REAL *D = M + ((size_t)k * 3 + 1) * ss;
        const REAL *R = M + ((size_t)k * 3 + 2) * ss;
        for (i = 0; i < s; i++) {
            REAL a0 = (REAL)0, a1 = (REAL)0, b0 = (REAL)0, b1 = (REAL)0, c0 = (REAL)0, c1 = (REAL)0;
            for (j = 0; j + 1 < s; j += 2) { a0 += D[i + j * s] * x[(size_t)k * s + j]; a1 += D[i + (j + 1) * s] * x[(size_t)k * s + j + 1]; }
            if (s & 1) a0 += D[i + (s - 1) * s] * x[(size_t)k * s + s - 1];
            if (k > 0) {
                for (j = 0; j + 1 < s; j += 2) { b0 += L[i + j * s] * x[(size_t)(k - 1) * s + j]; b1 += L[i + (j + 1) * s] * x[(size_t)(k - 1) * s + j + 1]; }
                if (s & 1) b0 += L[i + (s - 1) * s] * x[(size_t)(k - 1) * s + s - 1];
            }
            if (k < N - 1) {
                for (j = 0; j + 1 < s; j += 2) { c0 += R[i + j * s] * x[(size_t)(k + 1) * s + j]; c1 += R[i + (j + 1) * s] * x[(size_t)(k + 1) * s + j + 1]; }
                if (s & 1) c0 += R[i + (s - 1) * s] * x[(size_t)(k + 1) * s + s - 1];
            }
            y[(size_t)k * s + i] = ((b0 + b1) + (a0 + a1)) + (c0 + c1);
        }
    }
}

static REAL F(dot_alt)(int s, int N, const REAL *a, const REAL *b, REAL *work)
{
    int k, i, n = N;
    for (k = 0; k < N; k++) {
        REAL acc = (REAL)0;
        for (i = 0; i < s; i++) acc += a[(size_t)k * s + i] * b[(size_t)k * s + i];
        work[k] = acc;
    }
    while (n > 1) {
        int half = (n + 1) / 2;
        for (k = 0; k + half < n; k++) work[k] += work[k + half];
        n = half;
    }
    return work[0];
}

void F(oracle_pcg_mode)(int s, int N, const REAL *S, const REAL *Pinv, const REAL *gamma,
                        REAL *lambda, unsigned max_iter, REAL exit_tol,
                        unsigned *iters_out, unsigned char *max_iter_exit_out, int mode);

void F(oracle_pcg)(int s, int N, const REAL *S, const REAL *Pinv, const REAL *gamma,
                   REAL *lambda, unsigned max_iter, REAL exit_tol,
                   unsigned *iters_out, unsigned char *max_iter_exit_out)
{
    F(oracle_pcg_mode)(s, N, S, Pinv, gamma, lambda, max_iter, exit_tol, iters_out,
                       max_iter_exit_out, 0);
}

void F(oracle_pcg_mode)(int s, int N, const REAL *S, const REAL *Pinv, const REAL *gamma,
                        REAL *lambda, unsigned max_iter, REAL exit_tol,
                        unsigned *iters_out, unsigned char *max_iter_exit_out, int mode)
{
#define BDMV(M, x, y) do { if (mode) F(bdmv_alt)(s, N, M, x, y); else F(oracle_bdmv)(s, N, M, x, y); } while (0)
#define DOT(a, b) (mode ? F(dot_alt)(s, N, a, b, work) : F(dot)(n, a, b))
    const size_t n = (size_t)s * N;
    REAL *r = (REAL *)malloc(sizeof(REAL) * (n * 4 + N));
    REAL *p = r + n, *ups = p + n, *rt = ups + n, *work = rt + n;
    REAL eta, eta_new, alpha, beta, v;
    unsigned iter = 0;
    unsigned char max_iter_exit = 1;
    size_t i;

    BDMV(S, lambda, ups);
    for (i = 0; i < n; i++) r[i] = gamma[i] - ups[i];
    BDMV(Pinv, r, rt);
    for (i = 0; i < n; i++) p[i] = rt[i];
    eta = DOT(r, rt);

    /* upstream enters the loop unconditionally; an already-converged warm start
     * still costs one iteration unless eta is exactly 0 (guard against 0/0). */
    for (iter = 0; iter < max_iter; iter++) {
        BDMV(S, p, ups);
        v = DOT(p, ups);
        if (v == (REAL)0) { max_iter_exit = 0; break; }
        alpha = eta / v;
        for (i = 0; i < n; i++) { lambda[i] += alpha * p[i]; r[i] -= alpha * ups[i]; }
        BDMV(Pinv, r, rt);
        eta_new = DOT(r, rt);
        if (FABS(eta_new) < exit_tol) { iter++; max_iter_exit = 0; break; }
        beta = eta_new / eta;
        eta = eta_new;
        for (i = 0; i < n; i++) p[i] = rt[i] + beta * p[i];
    }
    *iters_out = iter;
    *max_iter_exit_out = max_iter_exit;
    free(r);
#undef BDMV
#undef DOT
}

/* ---------------------------------------------------------------------------
 * Single-reduction PCG (Chronopoulos & Gear 1989): the same Krylov iterates in exact
 * arithmetic, with the two inner products of an iteration (r.u and u.Su, u = Pinv r)
 * available together, so that a parallel implementation pays ONE global reduction per
 * iteration instead of two:   p.Sp = u.Su - beta^2 (p.Sp)_old.
 * Exit test, iteration count and max_iter flag as oracle_pcg.  This is the checker of the
 * product's MPCB200_PCG_VARIANT=1 builds; oracle_pcg stays the statement of the
 * reference's iteration.
 * ------------------------------------------------------------------------- */
void F(oracle_pcg_cg1)(int s, int N, const REAL *S, const REAL *Pinv, const REAL *gamma,
                       REAL *lambda, unsigned max_iter, REAL exit_tol,
                       unsigned *iters_out, unsigned char *max_iter_exit_out)
{
    const size_t n = (size_t)s * N;
    REAL *r = (REAL *)calloc(n * 5, sizeof(REAL));
    REAL *p = r + n, *q = p + n, *u = q + n, *w = u + n;     /* q = S p by recurrence */
    REAL gam, del, gam_old = (REAL)1, c_old = (REAL)0;       /* c = (p.Sp) / (r.u)^2 of the previous iteration */
    unsigned iter = 0;
    unsigned char max_iter_exit = 1;
    size_t i;

    F(oracle_bdmv)(s, N, S, lambda, w);
    for (i = 0; i < n; i++) r[i] = gamma[i] - w[i];
    F(oracle_bdmv)(s, N, Pinv, r, u);
    F(oracle_bdmv)(s, N, S, u, w);
    gam = F(dot)(n, r, u);
    del = F(dot)(n, u, w);
    for (iter = 0; iter < max_iter; iter++) {
        const REAL v = del - (gam * gam) * c_old;            /* p.Sp = u.Su - beta^2 (p.Sp)_old */
        const REAL beta = iter ? gam / gam_old : (REAL)0;
        REAL alpha;
        if (v == (REAL)0) { max_iter_exit = 0; break; }
        alpha = gam / v;
        for (i = 0; i < n; i++) {
            p[i] = u[i] + beta * p[i];
            q[i] = w[i] + beta * q[i];
            lambda[i] += alpha * p[i];
            r[i] -= alpha * q[i];
        }
        c_old = (v / gam) / gam;
        gam_old = gam;
        F(oracle_bdmv)(s, N, Pinv, r, u);
        F(oracle_bdmv)(s, N, S, u, w);
        gam = F(dot)(n, r, u);
        del = F(dot)(n, u, w);
        if (FABS(gam) < exit_tol) { iter++; max_iter_exit = 0; break; }
    }
    *iters_out = iter;
    *max_iter_exit_out = max_iter_exit;
    free(r);
}

/* ---------------------------------------------------------------------------
 * PCG with S p carried by a recurrence (product variant 2): the iteration of oracle_pcg with
 * ONE substitution -- after rt = Pinv r the product S rt is formed (on the GPU: while the
 * reduction of r.rt is in flight, which is the point) and
 *     p <- rt + beta p ,   S p <- S rt + beta (S p)
 * replaces the matvec S p of the next iteration.  alpha = eta / (p . Sp) and the exit test
 * are unchanged (two inner products per iteration, as the reference's iteration has).  Same
 * iterates in exact arithmetic; unlike the single-reduction form above no scalar is replaced
 * by a recurrence, only the vector S p.  This is the checker of the product's
 * MPCB200_PCG_VARIANT=2 builds (an A/B option, not the default); oracle_pcg stays the statement of the reference's
 * iteration and the parity criteria (oracle/parity.py) are asserted against THAT one.
 * ------------------------------------------------------------------------- */
void F(oracle_pcg_sprec)(int s, int N, const REAL *S, const REAL *Pinv, const REAL *gamma,
                         REAL *lambda, unsigned max_iter, REAL exit_tol,
                         unsigned *iters_out, unsigned char *max_iter_exit_out)
{
    const size_t n = (size_t)s * N;
    REAL *r = (REAL *)calloc(n * 5, sizeof(REAL));
    REAL *p = r + n, *q = p + n, *rt = q + n, *w = rt + n;   /* q = S p */
    REAL eta, eta_new, alpha, beta, v;
    unsigned iter = 0;
    unsigned char max_iter_exit = 1;
    size_t i;

    F(oracle_bdmv)(s, N, S, lambda, w);
    for (i = 0; i < n; i++) r[i] = gamma[i] - w[i];
    F(oracle_bdmv)(s, N, Pinv, r, rt);
    for (i = 0; i < n; i++) p[i] = rt[i];
    eta = F(dot)(n, r, rt);
    F(oracle_bdmv)(s, N, S, p, q);
    for (iter = 0; iter < max_iter; iter++) {
        v = F(dot)(n, p, q);
        if (v == (REAL)0) { max_iter_exit = 0; break; }
        alpha = eta / v;
        for (i = 0; i < n; i++) { lambda[i] += alpha * p[i]; r[i] -= alpha * q[i]; }
        F(oracle_bdmv)(s, N, Pinv, r, rt);
        F(oracle_bdmv)(s, N, S, rt, w);                      /* overlaps the reduction below */
        eta_new = F(dot)(n, r, rt);
        if (FABS(eta_new) < exit_tol) { iter++; max_iter_exit = 0; break; }
        beta = eta_new / eta;
        eta = eta_new;
        for (i = 0; i < n; i++) { p[i] = rt[i] + beta * p[i]; q[i] = w[i] + beta * q[i]; }
    }
    *iters_out = iter;
    *max_iter_exit_out = max_iter_exit;
    free(r);
}

/* ---------------------------------------------------------------------------
 * Values of the upper triangle of the symmetric block-tridiagonal S in CSC order,
 * restating store_block_csr_lowertri (include/utils/csr.cuh:10-36) as used by
 * form_schur_qdl_kernel (include/qdldl/linsys_setup.cuh:106,315,321): matrix column
 * j = block row b=j/s, local row=j%s holds [ S[b][0](row, 0..s-1) | S[b][1](row, 0..row) ].
 * nnz = (N-1)*s*s + N*s*(s+1)/2   (include/qdldl/sqp.cuh:148)
 * ------------------------------------------------------------------------- */
void F(oracle_csc_values)(int s, int N, const REAL *S, REAL *val)
{
    const int ss = s * s, tri = (s + 1) * s / 2, brow_ct = ss + tri;
    int b, row, col;
    for (b = 0; b < N; b++) {
        const REAL *L = S + ((size_t)b * 3 + 0) * ss;
        const REAL *D = S + ((size_t)b * 3 + 1) * ss;
        for (row = 0; row < s; row++) {
            size_t off = (size_t)(b > 0) * tri + (size_t)(b > 0) * (b - 1) * brow_ct
                       + (size_t)(b > 0) * row * s + (size_t)((row + 1) * row) / 2;
            if (b > 0) for (col = 0; col < s; col++) val[off + col] = L[row + col * s];
            for (col = 0; col <= row; col++) val[off + (b > 0) * s + col] = D[row + col * s];
        }
    }
}

/* ---------------------------------------------------------------------------
 * compute_dz: include/common/dz.cuh:3-121 (the consumer of
 * lambda and G^-1).  Ginv = G after oracle_schur, C and g as given to it.
 *   state row k   : dz = Qi_k ( q_k - (lambda_k + Abar_k^T lambda_{k+1}) )   (:59-113)
 *   control row k : dz = Ri_k ( r_k - Bbar_k^T lambda_{k+1} )               (:19-56)
 * dz: (s+c)*N - c.
 * ------------------------------------------------------------------------- */
void F(oracle_dz)(int s, int c, int N, const REAL *Ginv, const REAL *C, const REAL *g,
                  const REAL *lambda, REAL *dz)
{
    const int ss = s * s, cc = c * c, sc = s * c;
    const int Gset = ss + cc, Cset = ss + sc, gset = s + c;
    REAL scratch[ORACLE_MAX_DIM];
    int k, i, j;
    for (k = 0; k < N; k++) {
        const REAL *Qi = Ginv + (size_t)k * Gset;
        const REAL *lk = lambda + (size_t)k * s;
        for (i = 0; i < s; i++) {
            REAL acc = (REAL)0;
            if (k != N - 1) {
                const REAL *A = C + (size_t)k * Cset;
                for (j = 0; j < s; j++) acc += A[i * s + j] * lambda[(size_t)(k + 1) * s + j];
            }
            scratch[i] = lk[i] + acc;
            scratch[i] = g[(size_t)k * gset + i] - scratch[i];
        }
        F(matvec)(s, s, Qi, scratch, dz + (size_t)k * gset);
        if (k != N - 1) {
            const REAL *Ri = Qi + ss;
            const REAL *B = C + (size_t)k * Cset + ss;
            for (i = 0; i < c; i++) {
                REAL acc = (REAL)0;
                for (j = 0; j < s; j++) acc += B[i * s + j] * lambda[(size_t)(k + 1) * s + j];
                scratch[i] = g[(size_t)k * gset + s + i] - acc;
            }
            F(matvec)(c, c, Ri, scratch, dz + (size_t)k * gset + s);
        }
    }
}
