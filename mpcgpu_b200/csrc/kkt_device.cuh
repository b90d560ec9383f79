// kkt_device.cuh -- device-side building blocks:
//   schur_rows():  G,C,g,c,rho  ->  register-resident rows of S, Pinv, gamma (+ G^-1 to HBM)
//   pcg_rows():    register-resident S / Pinv rows -> lambda   (2 cluster barriers / iteration)
//
// Math restated from the reference (NOT a port of its kernels; different decomposition):
//   phase 1  include/pcg/linsys_setup.cuh:139-562   (S rows, Pinv diagonal, gamma, G^-1)
//   phase 2  include/pcg/linsys_setup.cuh:9-137     (SS preconditioner off-diagonals)
//   PCG      GBD-PCG gpu_pcg.cuh (absent; call site include/pcg/sqp.cuh:142-166,275)
//
// Decomposition: knot k's lane group inverts Qt_k and Rt_{k-1} ONCE (the reference inverts
// every Qt twice, in block rows k and k+1), forms phi_k, theta_k, gamma_k, theta_k^-1, and
// exchanges rows with knots k-1 / k+1 through shared memory (DSMEM across a cluster).
#pragma once
#include "common.cuh"

namespace mpcb200 {

template <int S_, int C_, int KPC_, int CL_>
struct Cfg {
    static constexpr int S = S_, C = C_, KPC = KPC_, CL = CL_;
    static constexpr int NT = KPC * LPK;        // threads per CTA
    static constexpr int NW = NT / 32;          // warps per CTA
    static constexpr int MT = S * LD;           // floats of one padded S x S tile
    static constexpr int BT = C * LD;           // padded B tile (C columns of S rows)
    static constexpr int RLD = 8;               // leading dim of the C x C tile (C <= 8)
    static constexpr int RT = C * RLD;
    static constexpr int SLOTS = KPC + 2;       // tiles with a halo slot on both sides
    static constexpr int NRED = CL * NW;        // partial sums per reduction
    static_assert(S <= 16 && C <= 8 && (KPC % 2) == 0, "unsupported block sizes");

    // shared-memory carve-up (float offsets)
    static constexpr int oM1 = 0;                       // [SLOTS][MT]  Qt^-1 rows, later Pinv-left rows
    static constexpr int oM2 = oM1 + SLOTS * MT;        // [SLOTS][MT]  phi rows
    static constexpr int oM3 = oM2 + SLOTS * MT;        // [SLOTS][MT]  theta^-1 rows
    static constexpr int oA = oM3 + SLOTS * MT;         // [KPC][MT]    A_{k-1} column-major padded
    static constexpr int oB = oA + KPC * MT;            // [KPC][BT]    B_{k-1} column-major padded
    static constexpr int oR = oB + KPC * BT;            // [KPC][RT]    Rt_{k-1}^-1 rows
    static constexpr int oV = oR + KPC * RT;            // [KPC][4][LD] q_k, q_{k-1}, r_{k-1}, c_k
    static constexpr int oPiv = oV + KPC * 4 * LD;      // [KPC][LD]    pivot-row scratch
    static constexpr int oPv = oPiv + KPC * LD;         // [SLOTS][LD]  p   (PCG)
    static constexpr int oRv = oPv + SLOTS * LD;        // [SLOTS][LD]  r   (PCG)
    static constexpr int oHal = oRv + SLOTS * LD;       // [2 bufs][2 sides][LD] pushed halo rows
    static constexpr int oRed = oHal + 4 * LD;          // [2 bufs][NRED (padded to 4)]
    static constexpr int NREDP = (NRED + 3) & ~3;
    static constexpr int FLOATS = oRed + 2 * NREDP;
    static constexpr size_t SMEM_BYTES = sizeof(float) * FLOATS;
};

// Per-instance views of the reference buffers (include/pcg/sqp.cuh:54-58).
struct KktPtrs {
    float* G;         // in: [Q_k | R_k]; out: [Qt_k^-1 | Rt_k^-1]
    const float* C;   // [Abar_k | Bbar_k]
    const float* g;   // [q_k | r_k]
    const float* c;   // constraint residuals
};

// Register-resident block row of the stored (negated) Schur system and preconditioner.
template <int S>
struct Rows {
    float Sl[S], Sd[S], Sr[S];
    float Pl[S], Pd[S], Pr[S];
    float gamma;
};

// copy one padded tile from a neighbour CTA's shared memory into a local halo slot
template <class CF>
__device__ __forceinline__ void pull_tile(float* smem, int off_tile_local, int off_tile_remote,
                                          unsigned src_rank) {
    const float4* src = reinterpret_cast<const float4*>(
        map_rank<CF::CL>(smem + off_tile_remote, src_rank));
    float4* dst = reinterpret_cast<float4*>(smem + off_tile_local);
    for (int q = threadIdx.x; q < CF::MT / 4; q += CF::NT) dst[q] = src[q];
}

// ---------------------------------------------------------------------------------------
// schur_rows: everything form_schur_system produces, for the knots owned by this CTA.
// Leaves S / Pinv rows and gamma_i in registers (for a fused solve) -- the two-stage
// kernel stores them to the reference layout afterwards.  Writes G^-1 in place.
// ---------------------------------------------------------------------------------------
template <class CF>
__device__ __forceinline__ void schur_rows(float* smem, const KktPtrs& P, int N, float rho,
                                           Rows<CF::S>& R) {
    constexpr int S = CF::S, C = CF::C, KPC = CF::KPC, CL = CF::CL, MT = CF::MT;
    constexpr int SS = S * S, CC = C * C, SC = S * C;
    constexpr int GSET = SS + CC, CSET = SS + SC, gSET = S + C;
    const int tid = threadIdx.x;
    const int g = tid / LPK, i = tid % LPK;
    const unsigned rank = cluster_rank<CL>();
    const int k0 = rank * KPC;         // first knot of this CTA
    const int k = k0 + g;
    const bool act = i < S;
    const bool valid = k < N;
    const bool inner = valid && k >= 1;   // block rows that have a left neighbour

    float* M1 = smem + CF::oM1;
    float* M2 = smem + CF::oM2;
    float* M3 = smem + CF::oM3;
    float* tA = smem + CF::oA + g * MT;
    float* tB = smem + CF::oB + g * CF::BT;
    float* tR = smem + CF::oR + g * CF::RT;
    float* vq = smem + CF::oV + g * 4 * LD;   // q_k
    float* vqm = vq + LD;                     // q_{k-1}
    float* vrm = vq + 2 * LD;                 // r_{k-1}
    float* vc = vq + 3 * LD;                  // c_k
    float* piv = smem + CF::oPiv + g * LD;

    // ---- loads ------------------------------------------------------------------------
    // [Abar_{k-1} | Bbar_{k-1}] is one contiguous CSET run per knot: CTA-cooperative,
    // coalesced copy into the padded column-major tiles.
    for (int idx = tid; idx < KPC * CSET; idx += CF::NT) {
        const int gg = idx / CSET, e = idx - gg * CSET;
        const int kk = k0 + gg;
        float v = 0.f;
        if (kk >= 1 && kk < N) v = __ldg(P.C + (size_t)(kk - 1) * CSET + e);
        if (e < SS) smem[CF::oA + gg * MT + (e / S) * LD + (e % S)] = v;
        else        smem[CF::oB + gg * CF::BT + ((e - SS) / S) * LD + ((e - SS) % S)] = v;
    }
    float aQ[S], aR[C];
#pragma unroll
    for (int j = 0; j < S; j++) {
        float v = 0.f;
        if (valid && act) v = P.G[(size_t)k * GSET + j * S + i];
        aQ[j] = v + ((j == i) ? rho : 0.f);                       // linsys_setup.cuh:180,329-330
    }
#pragma unroll
    for (int j = 0; j < C; j++) {
        float v = 0.f;
        if (inner && i < C) v = P.G[(size_t)(k - 1) * GSET + SS + j * C + i];
        aR[j] = v + ((j == i) ? rho : 0.f);                       // :331
    }
    {
        const bool a = act;
        vq[i] = (a && valid) ? __ldg(P.g + (size_t)k * gSET + i) : 0.f;
        vqm[i] = (a && inner) ? __ldg(P.g + (size_t)(k - 1) * gSET + i) : 0.f;
        vc[i] = (a && inner) ? __ldg(P.c + (size_t)k * S + i) : 0.f;
        if (i < CF::RLD) vrm[i] = (inner && i < C) ? __ldg(P.g + (size_t)(k - 1) * gSET + S + i) : 0.f;
    }
    // halo slots may be read (and discarded) by edge knots: keep them finite
    for (int q = tid; q < 2 * MT; q += CF::NT) {
        const int off = (q < MT) ? q : (KPC + 1) * MT + (q - MT);
        M1[off] = 0.f; M2[off] = 0.f; M3[off] = 0.f;
    }
    // knot 0: theta_0^-1 == Qt_0 (stored before inversion, linsys_setup.cuh:202-210)
    if (k == 0 && act) publish_row<S>(M3 + (g + 1) * MT, LD, i, aQ);

    // ---- phase 1a: inverses ---------------------------------------------------------------
    gj_invert_rows<S>(aQ, i, piv);                                 // Qt_k^-1
    gj_invert_rows<C>(aR, i, piv);                                 // Rt_{k-1}^-1
    if (valid && act) {
#pragma unroll
        for (int j = 0; j < S; j++) P.G[(size_t)k * GSET + j * S + i] = aQ[j];      // :372,:379
    }
    if (inner && i < C) {
#pragma unroll
        for (int j = 0; j < C; j++) P.G[(size_t)(k - 1) * GSET + SS + j * C + i] = aR[j];  // :375
    }
    if (act) publish_row<S>(M1 + (g + 1) * MT, LD, i, aQ);
    if (i < C) publish_row<C>(tR, CF::RLD, i, aR);
    __syncwarp();
    const float t_i = row_dot<S>(aQ, vq);                          // Qt_k^-1 q_k  (:411-415, :260)

    cluster_sync_all<CL>();
    if (CL > 1 && rank > 0) pull_tile<CF>(smem, CF::oM1, CF::oM1 + KPC * MT, rank - 1);
    if (CL > 1) __syncthreads();

    // ---- phase 1b: phi, theta, gamma, theta^-1 ---------------------------------------------
    float th[S];
    {
        float phi[S], br[C], arow[S], brow[C];
#pragma unroll
        for (int m = 0; m < S; m++) arow[m] = tA[m * LD + i];      // A_{k-1}[i][m]
#pragma unroll
        for (int m = 0; m < C; m++) brow[m] = tB[m * LD + i];      // B_{k-1}[i][m]
#pragma unroll
        for (int j = 0; j < S; j++) phi[j] = 0.f;
#pragma unroll
        for (int j = 0; j < C; j++) br[j] = 0.f;
        row_gemm_acc<S, S>(phi, arow, M1 + g * MT, LD);            // phi = A Qt_{k-1}^-1   (:398)
        row_gemm_acc<C, C>(br, brow, tR, CF::RLD);                 // BR  = B Rt_{k-1}^-1   (:406)
#pragma unroll
        for (int j = 0; j < S; j++) th[j] = 0.f;
        row_gemm_acc<S, S>(th, phi, tA, LD);                       // phi A^T               (:447)
#pragma unroll
        for (int j = 0; j < S; j++) th[j] += aQ[j];                // + Qt_k^-1             (:466)
        {
            float t2[S];
#pragma unroll
            for (int j = 0; j < S; j++) t2[j] = 0.f;
            row_gemm_acc<C, S>(t2, br, tB, LD);                    // BR B^T                (:473)
#pragma unroll
            for (int j = 0; j < S; j++) th[j] += t2[j];            //                       (:485)
        }
        const float gam = (t_i - vc[i]) + (row_dot<C>(br, vrm) + row_dot<S>(phi, vqm));  // :416-443
        if (inner) {
#pragma unroll
            for (int j = 0; j < S; j++) { R.Sl[j] = -phi[j]; R.Sd[j] = -th[j]; }   // :491,:501
            R.gamma = -gam;                                                         // :531
        } else {
#pragma unroll
            for (int j = 0; j < S; j++) { R.Sl[j] = 0.f; R.Sd[j] = -aQ[j]; }       // :249
            R.gamma = -t_i;                                                         // :274
        }
        if (act) publish_row<S>(M2 + (g + 1) * MT, LD, i, phi);
    }
    float thi[S];
#pragma unroll
    for (int j = 0; j < S; j++) thi[j] = th[j];
    gj_invert_rows<S>(thi, i, piv);                                // theta_k^-1            (:511-513)
    if (k != 0 && act) publish_row<S>(M3 + (g + 1) * MT, LD, i, thi);
    __syncwarp();
    if (k == 0) load_row<S>(thi, M3 + (g + 1) * MT + (act ? i : 0) * LD);   // Qt_0 row
#pragma unroll
    for (int j = 0; j < S; j++) R.Pd[j] = -thi[j];                 // :518-524 / :202-210

    cluster_sync_all<CL>();
    if (CL > 1) {
        if (rank > 0) pull_tile<CF>(smem, CF::oM3, CF::oM3 + KPC * MT, rank - 1);
        if (rank + 1 < (unsigned)CL)
            pull_tile<CF>(smem, CF::oM2 + (KPC + 1) * MT, CF::oM2 + 1 * MT, rank + 1);
        __syncthreads();
    }

    // ---- phase 2: S right block, Pinv left block ---------------------------------------------
    const bool has_right = valid && (k + 1 < N);
    {
        const float* phin = M2 + (g + 2) * MT;                     // phi_{k+1} rows
#pragma unroll
        for (int j = 0; j < S; j++) R.Sr[j] = (has_right && act) ? -phin[j * LD + i] : 0.f;  // :537-557
    }
    {
        float t1[S], pl[S];
#pragma unroll
        for (int j = 0; j < S; j++) { t1[j] = 0.f; pl[j] = 0.f; }
        row_gemm_acc<S, S>(t1, thi, M2 + (g + 1) * MT, LD);        // theta_k^-1 phi_k        (:100)
        row_gemm_acc<S, S>(pl, t1, M3 + g * MT, LD);               // ... theta_{k-1}^-1      (:102)
#pragma unroll
        for (int j = 0; j < S; j++) R.Pl[j] = inner ? pl[j] : 0.f; // sign: see DESIGN.md
    }
    // M1 (Qt^-1 rows) is dead after phase 1b on every CTA (barrier above): reuse for Pinv-left
    if (act) publish_row<S>(M1 + (g + 1) * MT, LD, i, R.Pl);

    cluster_sync_all<CL>();
    if (CL > 1) {
        if (rank + 1 < (unsigned)CL)
            pull_tile<CF>(smem, CF::oM1 + (KPC + 1) * MT, CF::oM1 + 1 * MT, rank + 1);
        __syncthreads();
    }
    {
        const float* pln = M1 + (g + 2) * MT;                      // Pinv_left_{k+1} rows
#pragma unroll
        for (int j = 0; j < S; j++) R.Pr[j] = (has_right && act) ? pln[j * LD + i] : 0.f;   // :118-136
    }
    if (!valid || !act) {
#pragma unroll
        for (int j = 0; j < S; j++) { R.Sl[j] = R.Sd[j] = R.Sr[j] = 0.f; R.Pl[j] = R.Pd[j] = R.Pr[j] = 0.f; }
        R.gamma = 0.f;
    }
}

// ---------------------------------------------------------------------------------------
// Cluster-wide sum with a bit-identical result in every thread of every CTA: warp shuffle,
// one DSMEM store per (warp, destination CTA), ONE cluster barrier, then every warp
// re-reduces the same NRED partials with the same tree.
// ---------------------------------------------------------------------------------------
template <class CF>
__device__ __forceinline__ float cluster_sum(float* red, int buf, float val, unsigned rank) {
    constexpr int CL = CF::CL, NW = CF::NW, NRED = CF::NRED;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float w = warp_sum(val);
    float* slot = red + buf * CF::NREDP + rank * NW + warp;
    if constexpr (CL == 1) {
        if (lane == 0) *slot = w;
    } else {
        if (lane < CL) *map_rank<CL>(slot, (unsigned)lane) = w;
    }
    cluster_sync_all<CL>();
    float acc = 0.f;
#pragma unroll
    for (int q = 0; q < (NRED + 31) / 32; q++) {
        const int idx = lane + 32 * q;
        if (idx < NRED) acc += red[buf * CF::NREDP + idx];
    }
    return warp_sum(acc);
}

template <int S>
__device__ __forceinline__ float bd_row(const float (&L)[S], const float (&D)[S],
                                        const float (&Rr)[S], const float* vec, int g) {
    return row_dot<S>(L, vec + g * LD) + row_dot<S>(D, vec + (g + 1) * LD) +
           row_dot<S>(Rr, vec + (g + 2) * LD);
}

// ---------------------------------------------------------------------------------------
// pcg_rows: preconditioned CG on the stored system S lambda = gamma, S / Pinv rows in
// registers.  Semantics (warm start, |r.Pinv r| < tol absolute exit, iteration count,
// max-iter flag) follow SURVEY 0.2 / oracle_pcg.  Per iteration: 2 cluster barriers (one per
// dot product) + 2 CTA barriers; the boundary rows of S p and Pinv r ride along with the
// partial sums, so neighbours update their halo copies of r and p locally.
// ---------------------------------------------------------------------------------------
template <class CF>
__device__ __forceinline__ void pcg_rows(float* smem, const Rows<CF::S>& R, float* lam_g, int N,
                                         unsigned max_iter, float exit_tol, unsigned* d_iters,
                                         uint8_t* d_exit) {
    constexpr int S = CF::S, KPC = CF::KPC, CL = CF::CL;
    const int tid = threadIdx.x;
    const int g = tid / LPK, i = tid % LPK;
    const unsigned rank = cluster_rank<CL>();
    const int k0 = rank * KPC, k = k0 + g;
    const bool act = i < S, valid = k < N;
    const bool first_g = (g == 0), last_g = (g == KPC - 1);
    const bool has_l = rank > 0, has_r = rank + 1 < (unsigned)CL;

    float* pv = smem + CF::oPv;
    float* rv = smem + CF::oRv;
    float* hal = smem + CF::oHal;   // [buf][side][LD]; side 0: row of my LEFT neighbour knot
    float* red = smem + CF::oRed;
    float* my_p = pv + (g + 1) * LD + i;
    float* my_r = rv + (g + 1) * LD + i;
    float* hl_p = pv + i;                    // left halo slot, lane i
    float* hr_p = pv + (KPC + 1) * LD + i;   // right halo slot
    float* hl_r = rv + i;
    float* hr_r = rv + (KPC + 1) * LD + i;

    // ---- r = gamma - S lambda0 --------------------------------------------------------------
    float lam_i = (valid && act) ? lam_g[(size_t)k * S + i] : 0.f;
    *my_p = lam_i;
    if (first_g) *hl_p = (k0 >= 1 && k0 - 1 < N && act) ? lam_g[(size_t)(k0 - 1) * S + i] : 0.f;
    if (last_g) *hr_p = (k0 + KPC < N && act) ? lam_g[(size_t)(k0 + KPC) * S + i] : 0.f;
    __syncthreads();
    float r_i = R.gamma - bd_row<S>(R.Sl, R.Sd, R.Sr, pv, g);
    *my_r = r_i;
    if (first_g) { if (has_l) *map_rank<CL>(hr_r, rank - 1) = r_i; else *hl_r = 0.f; }
    if (last_g)  { if (has_r) *map_rank<CL>(hl_r, rank + 1) = r_i; else *hr_r = 0.f; }
    cluster_sync_all<CL>();

    // ---- p = rt = Pinv r ; eta = r.rt ----------------------------------------------------------
    float rt = bd_row<S>(R.Pl, R.Pd, R.Pr, rv, g);
    float p_i = rt;
    *my_p = p_i;       // pv was last read before the barrier above
    if (first_g) { if (has_l) *map_rank<CL>(hr_p, rank - 1) = p_i; else *hl_p = 0.f; }
    if (last_g)  { if (has_r) *map_rank<CL>(hl_p, rank + 1) = p_i; else *hr_p = 0.f; }
    float eta = cluster_sum<CF>(red, 1, r_i * rt, rank);

    unsigned iter = 0;
    bool max_iter_exit = true;
    for (; iter < max_iter; iter++) {
        // ---- ups = S p ; v = p.ups -----------------------------------------------------------
        const float ups = bd_row<S>(R.Sl, R.Sd, R.Sr, pv, g);
        if (CL > 1) {
            if (first_g && has_l) *map_rank<CL>(hal + (0 * 2 + 1) * LD + i, rank - 1) = ups;
            if (last_g && has_r)  *map_rank<CL>(hal + (0 * 2 + 0) * LD + i, rank + 1) = ups;
        }
        const float v = cluster_sum<CF>(red, 0, p_i * ups, rank);
        if (v == 0.f) { max_iter_exit = false; break; }
        const float alpha = eta / v;
        lam_i = fmaf(alpha, p_i, lam_i);
        r_i = fmaf(-alpha, ups, r_i);
        *my_r = r_i;
        if (CL > 1) {
            if (first_g && has_l) *hl_r = fmaf(-alpha, hal[(0 * 2 + 0) * LD + i], *hl_r);
            if (last_g && has_r)  *hr_r = fmaf(-alpha, hal[(0 * 2 + 1) * LD + i], *hr_r);
        }
        __syncthreads();
        // ---- rt = Pinv r ; eta_new = r.rt ----------------------------------------------------
        rt = bd_row<S>(R.Pl, R.Pd, R.Pr, rv, g);
        if (CL > 1) {
            if (first_g && has_l) *map_rank<CL>(hal + (1 * 2 + 1) * LD + i, rank - 1) = rt;
            if (last_g && has_r)  *map_rank<CL>(hal + (1 * 2 + 0) * LD + i, rank + 1) = rt;
        }
        const float eta_new = cluster_sum<CF>(red, 1, r_i * rt, rank);
        if (fabsf(eta_new) < exit_tol) { iter++; max_iter_exit = false; break; }
        const float beta = eta_new / eta;
        eta = eta_new;
        p_i = fmaf(beta, p_i, rt);
        *my_p = p_i;
        if (CL > 1) {
            if (first_g && has_l) *hl_p = fmaf(beta, *hl_p, hal[(1 * 2 + 0) * LD + i]);
            if (last_g && has_r)  *hr_p = fmaf(beta, *hr_p, hal[(1 * 2 + 1) * LD + i]);
        }
        __syncthreads();
    }
    if (valid && act) lam_g[(size_t)k * S + i] = lam_i;
    if (rank == 0 && tid == 0) { *d_iters = iter; *d_exit = max_iter_exit ? 1 : 0; }
    // no CTA may exit while a peer can still address its shared memory
    if (CL > 1) cluster_sync_all<CL>();
}

}  // namespace mpcb200
