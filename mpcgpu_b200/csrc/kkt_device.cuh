// kkt_device.cuh -- device-side building blocks of the sm_100a KKT linear-system kernels:
//   schur_rows():  G,C,g,c,rho -> S rows in registers, Pinv rows in shared memory, gamma
//                  (+ G^-1 written in place to HBM)
//   pcg_rows():    S rows (registers) + Pinv rows (shared memory) -> lambda
//
// Math restated from the reference (NOT a port of its kernels; different decomposition):
//   phase 1  include/pcg/linsys_setup.cuh:139-562   (S rows, Pinv diagonal, gamma, G^-1)
//   phase 2  include/pcg/linsys_setup.cuh:9-137     (SS preconditioner off-diagonals)
//   PCG      GBD-PCG gpu_pcg.cuh (absent; call site include/pcg/sqp.cuh:142-166,275)
//
// Mapping ("row-pair owner"): 8 lanes per knot, lane t < 7 owns rows t and t+7 of every
// 14 x 14 block of its knot (lane 7 idles: 87.5 % lane use, the same as 14 of 16).  Four knots
// share a warp, so everything inside a knot synchronises with __syncwarp().  One CTA holds up
// to 64 knots (512 threads); a whole 64-knot instance lives in ONE SM:
//   * S block rows (3 x 14 floats per row) stay in registers for the whole solve (84 regs);
//   * Pinv block rows live in shared memory as PROW[knot][row][44] = [Pd | Pl | Pr | pad2]
//     (row stride 44 floats: 16 B aligned and bank-conflict free for the 7 active lanes);
//   * all pair-wise arithmetic is FFMA2 (fma.rn.f32x2, sm_100): half the FMA issue slots.
// The Schur phase uses the same shared memory as scratch: the Pd columns of PROW first hold
// Qt^-1 rows, then -theta^-1 rows; the (not yet written) Pl / Pr columns hold B, R^-1 and
// the knot's vectors; a second tile per knot holds A, then phi.
// Longer horizons use a thread-block cluster of CL CTAs (64 knots each) with DSMEM halos.
#pragma once
#include "common.cuh"

namespace mpcb200 {

template <int KPC_, int CL_>
struct Cfg {
    static constexpr int S = 14, C = 7, H = 7;
    static constexpr int KPC = KPC_, CL = CL_;
    static constexpr int LPK = 8;               // lanes per knot
    static constexpr int NT = KPC * LPK;        // threads per CTA
    static constexpr int NW = NT / 32;          // warps per CTA
    static constexpr int SLOTS = KPC + 2;       // one halo knot on both sides
    static constexpr int PR = 44;               // floats per PROW row
    static constexpr int PK = S * PR;           // floats per PROW knot (616)
    static constexpr int TK = S * S;            // floats per A / phi tile (196)
    static constexpr int VK = 16;               // floats per vector slot
    static constexpr int cD = 0, cL = 14, cR = 28;   // column offsets of Pd | Pl | Pr in a row
    // shared-memory carve-up (float offsets)
    static constexpr int oP = 0;                        // [SLOTS][14][44]  PROW
    static constexpr int oT = oP + SLOTS * PK;          // [SLOTS][196]     A (col-major) -> phi (row-major)
    static constexpr int oPv = oT + SLOTS * TK;         // [SLOTS][16]      p
    static constexpr int oRv = oPv + SLOTS * VK;        // [SLOTS][16]      r
    static constexpr int oHal = oRv + SLOTS * VK;       // [2 phases][2 sides][16] pushed boundary rows
    static constexpr int NRED = CL * NW;                // partial sums per reduction
    static constexpr int NREDP = (NRED + 3) & ~3;
    static constexpr int oRed = oHal + 4 * VK;          // [2 bufs][NREDP]
    static constexpr int FLOATS = oRed + 2 * NREDP;
    static constexpr size_t SMEM_BYTES = sizeof(float) * FLOATS;
    static_assert(KPC % 4 == 0 && KPC <= 128, "whole warps, <= 1024 threads");
    static_assert(SMEM_BYTES <= 232448, "exceeds the 227 KB per-CTA shared memory of sm_100");
};

// Per-instance views of the reference buffers (include/pcg/sqp.cuh:54-58).
struct KktPtrs {
    float* G;         // in: [Q_k | R_k]; out: [Qt_k^-1 | Rt_k^-1]
    const float* C;   // [Abar_k | Bbar_k]
    const float* g;   // [q_k | r_k]
    const float* c;   // constraint residuals
};

// Register-resident part of a knot's block row pair: rows t and t+7 of the stored (negated)
// Schur complement, as float2 pairs.
struct Rows {
    float2 Sl[2][7], Sd[2][7], Sr[2][7];
    float gamma[2];
};

// ---- small float2 helpers ----------------------------------------------------------------
__device__ __forceinline__ float2 f2(float a, float b) { return make_float2(a, b); }
__device__ __forceinline__ float2 splat(float a) { return make_float2(a, a); }
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ float getp(const float2 (&a)[7], int i) { return (i & 1) ? a[i >> 1].y : a[i >> 1].x; }
__device__ __forceinline__ void setp(float2 (&a)[7], int i, float v) { if (i & 1) a[i >> 1].y = v; else a[i >> 1].x = v; }

// 14 floats from a 16 B aligned address (reads 16 floats; the last two are ignored)
__device__ __forceinline__ void ld14_v4(float2 (&r)[7], const float* p) {
    const float4* s = reinterpret_cast<const float4*>(p);
    const float4 a = s[0], b = s[1], c = s[2], d = s[3];
    r[0] = f2(a.x, a.y); r[1] = f2(a.z, a.w); r[2] = f2(b.x, b.y); r[3] = f2(b.z, b.w);
    r[4] = f2(c.x, c.y); r[5] = f2(c.z, c.w); r[6] = f2(d.x, d.y);
}
// 14 floats from an 8 B aligned address
__device__ __forceinline__ void ld14_v2(float2 (&r)[7], const float* p) {
    const float2* s = reinterpret_cast<const float2*>(p);
#pragma unroll
    for (int q = 0; q < 7; q++) r[q] = s[q];
}
// 14 floats to a 16 B aligned address (exactly 14 floats are written)
__device__ __forceinline__ void st14_v4(float* p, const float2 (&r)[7]) {
    float4* d = reinterpret_cast<float4*>(p);
    d[0] = make_float4(r[0].x, r[0].y, r[1].x, r[1].y);
    d[1] = make_float4(r[2].x, r[2].y, r[3].x, r[3].y);
    d[2] = make_float4(r[4].x, r[4].y, r[5].x, r[5].y);
    reinterpret_cast<float2*>(p)[6] = r[6];
}
__device__ __forceinline__ void st14_v2(float* p, const float2 (&r)[7]) {
    float2* d = reinterpret_cast<float2*>(p);
#pragma unroll
    for (int q = 0; q < 7; q++) d[q] = r[q];
}

// dot of a 14-row (float2 pairs) with 14 floats: two interleaved partial sums (even / odd)
__device__ __forceinline__ float2 dot14_acc(const float2 (&a)[7], const float2 (&x)[7], float2 acc) {
#pragma unroll
    for (int q = 0; q < 7; q++) acc = fma2(a[q], x[q], acc);
    return acc;
}

// c0 += a0 * b, c1 += a1 * b  (one broadcast row b shared by the thread's two rows)
__device__ __forceinline__ void axpy2(float2 (&c0)[7], float2 (&c1)[7], float a0, float a1,
                                      const float2 (&b)[7]) {
    const float2 s0 = splat(a0), s1 = splat(a1);
#pragma unroll
    for (int q = 0; q < 7; q++) { c0[q] = fma2(s0, b[q], c0[q]); c1[q] = fma2(s1, b[q], c1[q]); }
}

// In-place Gauss-Jordan inverse WITHOUT pivoting of a 14 x 14 matrix held as row pairs (lane t
// owns rows t and t+7).  Same elimination sequence as the reference's invertMatrix on [A | I]
// (include/utils/matrix.cuh:120-238): the in-place form keeps the evolving inverse column in
// the slot of the eliminated one.  `piv`: 16 floats of shared memory private to this knot.
__device__ __forceinline__ void gj14(float2 (&a0)[7], float2 (&a1)[7], int t, bool act, float* piv) {
#pragma unroll
    for (int p = 0; p < 14; p++) {
        const int owner = p % 7;
        const bool hi = p >= 7;
        if (act && t == owner) { if (hi) st14_v4(piv, a1); else st14_v4(piv, a0); }
        __syncwarp();
        float2 rp[7];
        ld14_v4(rp, piv);
        __syncwarp();
        const float pinv = 1.0f / getp(rp, p);
        const bool own0 = !hi && t == owner, own1 = hi && t == owner;
        {
            const float f = getp(a0, p) * pinv;
            if (own0) {
#pragma unroll
                for (int q = 0; q < 7; q++) a0[q] = f2(rp[q].x * pinv, rp[q].y * pinv);
                setp(a0, p, pinv);
            } else {
                const float2 nf = splat(-f);
#pragma unroll
                for (int q = 0; q < 7; q++) a0[q] = fma2(nf, rp[q], a0[q]);
                setp(a0, p, -f);
            }
        }
        {
            const float f = getp(a1, p) * pinv;
            if (own1) {
#pragma unroll
                for (int q = 0; q < 7; q++) a1[q] = f2(rp[q].x * pinv, rp[q].y * pinv);
                setp(a1, p, pinv);
            } else {
                const float2 nf = splat(-f);
#pragma unroll
                for (int q = 0; q < 7; q++) a1[q] = fma2(nf, rp[q], a1[q]);
                setp(a1, p, -f);
            }
        }
    }
}

// 7 x 7 Gauss-Jordan, lane t < 7 owns row t (8 floats, last one padding).  piv: 8 floats.
__device__ __forceinline__ void gj7(float (&a)[8], int t, bool act, float* piv) {
#pragma unroll
    for (int p = 0; p < 7; p++) {
        if (act && t == p) {
            reinterpret_cast<float4*>(piv)[0] = make_float4(a[0], a[1], a[2], a[3]);
            reinterpret_cast<float4*>(piv)[1] = make_float4(a[4], a[5], a[6], 0.f);
        }
        __syncwarp();
        const float4 u = reinterpret_cast<const float4*>(piv)[0], v = reinterpret_cast<const float4*>(piv)[1];
        __syncwarp();
        const float rp[7] = {u.x, u.y, u.z, u.w, v.x, v.y, v.z};
        const float pinv = 1.0f / rp[p];
        const float f = a[p] * pinv;
        if (t == p) {
#pragma unroll
            for (int j = 0; j < 7; j++) a[j] = rp[j] * pinv;
            a[p] = pinv;
        } else {
#pragma unroll
            for (int j = 0; j < 7; j++)
                if (j != p) a[j] = fmaf(-f, rp[j], a[j]);
            a[p] = -f;
        }
    }
}

// copy one neighbour-CTA region into local shared memory (cluster kernels only)
template <class CF>
__device__ __forceinline__ void pull_floats(float* dst, const float* src_remote_local_addr, int n,
                                            unsigned src_rank) {
    const float2* src = reinterpret_cast<const float2*>(map_rank<CF::CL>(src_remote_local_addr, src_rank));
    float2* d = reinterpret_cast<float2*>(dst);
    for (int q = threadIdx.x; q < n / 2; q += CF::NT) d[q] = src[q];
}
// Pd columns (14 x 14, row stride PR) of one PROW knot from a neighbour CTA
template <class CF>
__device__ __forceinline__ void pull_pd(float* smem, int dst_slot, int src_slot, unsigned src_rank) {
    const float* src = map_rank<CF::CL>(smem + CF::oP + src_slot * CF::PK, src_rank);
    float* dst = smem + CF::oP + dst_slot * CF::PK;
    for (int q = threadIdx.x; q < 14 * 7; q += CF::NT) {
        const int row = q / 7, c2 = q - row * 7;
        reinterpret_cast<float2*>(dst + row * CF::PR)[c2] = reinterpret_cast<const float2*>(src + row * CF::PR)[c2];
    }
}

// ---------------------------------------------------------------------------------------
// schur_rows: everything form_schur_system produces, for the knots owned by this CTA.
// Leaves S rows and gamma in registers and the Pinv rows in PROW; writes G^-1 in place.
// ---------------------------------------------------------------------------------------
template <class CF>
__device__ __forceinline__ void schur_rows(float* smem, const KktPtrs& P, int N, float rho, Rows& R) {
    constexpr int S = 14, C = 7, KPC = CF::KPC, CL = CF::CL, PR = CF::PR, PK = CF::PK, TK = CF::TK;
    constexpr int SS = S * S, CC = C * C, SC = S * C;
    constexpr int GSET = SS + CC, CSET = SS + SC, gSET = S + C;
    constexpr int cD = CF::cD, cL = CF::cL, cR = CF::cR;
    const int tid = threadIdx.x;
    const int g = tid >> 3, t = tid & 7;
    const unsigned rank = cluster_rank<CL>();
    const int k0 = rank * KPC;
    const int k = k0 + g;
    const bool act = t < 7;
    const int r0 = act ? t : 0, r1 = r0 + 7;      // idle lane shadows rows 0 / 7 (never writes)
    const bool valid = k < N;
    const bool inner = valid && k >= 1;
    const bool has_right = valid && (k + 1 < N);
    const int sl = g + 1;

    float* PW = smem + CF::oP;                    // PROW
    float* TA = smem + CF::oT;
    float* myP = PW + sl * PK;
    float* myT = TA + sl * TK;
    // scratch living in the not-yet-written Pl / Pr columns of this knot's PROW rows
    float* vq = myP + 0 * PR + cR;                // q_k
    float* vqm = myP + 1 * PR + cR;               // q_{k-1}
    float* vc = myP + 2 * PR + cR;                // c_k
    float* vrm = myP + 3 * PR + cR;               // r_{k-1} (7)
    float* piv = myP + 4 * PR + cR;               // pivot row broadcast (16 B aligned, + 2 pad)
    // B_{k-1} column m: myP[m*PR + cL .. +13] (m < 7); Rt^-1 row m: myP[(7+m)*PR + cL .. +7]

    // ---- stage [Abar_{k-1} | Bbar_{k-1}] (one contiguous run per knot), coalesced ----------
    for (int idx = tid; idx < KPC * CSET; idx += CF::NT) {
        const int gg = idx / CSET, e = idx - gg * CSET;
        const int kk = k0 + gg;
        float v = 0.f;
        if (kk >= 1 && kk < N) v = __ldg(P.C + (size_t)(kk - 1) * CSET + e);
        if (e < SS) TA[(gg + 1) * TK + e] = v;                                   // column-major A
        else { const int m = (e - SS) / S, i = (e - SS) - m * S; PW[(gg + 1) * PK + m * PR + cL + i] = v; }
    }
    // halo knots: keep every tile that an edge knot may read finite
    for (int q = tid; q < 2 * PK; q += CF::NT) PW[(q < PK) ? q : (KPC + 1) * PK + (q - PK)] = 0.f;
    for (int q = tid; q < 2 * TK; q += CF::NT) TA[(q < TK) ? q : (KPC + 1) * TK + (q - TK)] = 0.f;

    // ---- loads: Qt_k rows, Rt_{k-1} row, vectors -----------------------------------------
    float2 aQ0[7], aQ1[7];
#pragma unroll
    for (int j = 0; j < 14; j++) {
        float v0 = 0.f, v1 = 0.f;
        if (valid && act) { v0 = P.G[(size_t)k * GSET + j * S + r0]; v1 = P.G[(size_t)k * GSET + j * S + r1]; }
        setp(aQ0, j, v0 + ((j == r0) ? rho : 0.f));                    // linsys_setup.cuh:180,329-330
        setp(aQ1, j, v1 + ((j == r1) ? rho : 0.f));
    }
    float aR[8];
#pragma unroll
    for (int j = 0; j < 7; j++) {
        float v = 0.f;
        if (inner && act) v = P.G[(size_t)(k - 1) * GSET + SS + j * C + r0];
        aR[j] = v + ((j == r0) ? rho : 0.f);                           // :331
    }
    aR[7] = 0.f;
    __syncthreads();                               // staging done (vectors overwrite zeroed chunks)
    if (act) {
        vq[r0] = valid ? __ldg(P.g + (size_t)k * gSET + r0) : 0.f;
        vq[r1] = valid ? __ldg(P.g + (size_t)k * gSET + r1) : 0.f;
        vqm[r0] = inner ? __ldg(P.g + (size_t)(k - 1) * gSET + r0) : 0.f;
        vqm[r1] = inner ? __ldg(P.g + (size_t)(k - 1) * gSET + r1) : 0.f;
        vc[r0] = inner ? __ldg(P.c + (size_t)k * S + r0) : 0.f;
        vc[r1] = inner ? __ldg(P.c + (size_t)k * S + r1) : 0.f;
        vrm[r0] = inner ? __ldg(P.g + (size_t)(k - 1) * gSET + S + r0) : 0.f;
    }
    if (t == 7) vrm[7] = 0.f;
    // knot 0: Pinv[0][1] = -Qt_0, stored BEFORE inversion (linsys_setup.cuh:202-210); its A
    // tile is unused, park the rows there
    if (k == 0 && act) { st14_v2(myT + r0 * S, aQ0); st14_v2(myT + r1 * S, aQ1); }

    // ---- phase 1a: inverses ---------------------------------------------------------------
    gj14(aQ0, aQ1, t, act, piv);                                       // Qt_k^-1
    gj7(aR, t, act, piv);                                              // Rt_{k-1}^-1
    if (valid && act) {
#pragma unroll
        for (int j = 0; j < 14; j++) {                                 // :372,:379
            P.G[(size_t)k * GSET + j * S + r0] = getp(aQ0, j);
            P.G[(size_t)k * GSET + j * S + r1] = getp(aQ1, j);
        }
    }
    if (inner && act) {
#pragma unroll
        for (int j = 0; j < 7; j++) P.G[(size_t)(k - 1) * GSET + SS + j * C + r0] = aR[j];   // :375
    }
    if (act) {
        st14_v4(myP + r0 * PR + cD, aQ0);
        st14_v4(myP + r1 * PR + cD, aQ1);
        float2* rr = reinterpret_cast<float2*>(myP + (7 + r0) * PR + cL);
        rr[0] = f2(aR[0], aR[1]); rr[1] = f2(aR[2], aR[3]); rr[2] = f2(aR[4], aR[5]); rr[3] = f2(aR[6], 0.f);
    }
    __syncwarp();
    float ti0, ti1;                                                    // Qt_k^-1 q_k  (:411-415, :260)
    {
        float2 x[7];
        ld14_v4(x, vq);
        const float2 s0 = dot14_acc(aQ0, x, splat(0.f)), s1 = dot14_acc(aQ1, x, splat(0.f));
        ti0 = s0.x + s0.y; ti1 = s1.x + s1.y;
    }

    cluster_sync_all<CL>();
    if (CL > 1) {
        if (rank > 0) pull_pd<CF>(smem, 0, KPC, rank - 1);
        __syncthreads();
    }

    // ---- phase 1b: phi, theta, gamma, theta^-1 ----------------------------------------------
    float2 th0[7], th1[7];
    {
        float2 phi0[7], phi1[7];
#pragma unroll
        for (int q = 0; q < 7; q++) { phi0[q] = splat(0.f); phi1[q] = splat(0.f); th0[q] = aQ0[q]; th1[q] = aQ1[q]; }   // + Qt_k^-1 (:466)
        const float* Qm = PW + (sl - 1) * PK + cD;                     // Qt_{k-1}^-1 rows
#pragma unroll
        for (int m = 0; m < 14; m++) {                                 // phi = A Qt_{k-1}^-1   (:398)
            float2 b[7];
            ld14_v4(b, Qm + m * PR);
            axpy2(phi0, phi1, myT[m * S + r0], myT[m * S + r1], b);
        }
        float br0[7], br1[7];
#pragma unroll
        for (int j = 0; j < 7; j++) { br0[j] = 0.f; br1[j] = 0.f; }
#pragma unroll
        for (int m = 0; m < 7; m++) {                                  // BR = B Rt_{k-1}^-1    (:406)
            const float b0 = myP[m * PR + cL + r0], b1 = myP[m * PR + cL + r1];
            const float2* rr = reinterpret_cast<const float2*>(myP + (7 + m) * PR + cL);
            const float2 u0 = rr[0], u1 = rr[1], u2 = rr[2], u3 = rr[3];
            const float rm[7] = {u0.x, u0.y, u1.x, u1.y, u2.x, u2.y, u3.x};
#pragma unroll
            for (int j = 0; j < 7; j++) { br0[j] = fmaf(b0, rm[j], br0[j]); br1[j] = fmaf(b1, rm[j], br1[j]); }
        }
#pragma unroll
        for (int m = 0; m < 14; m++) {                                 // + phi A^T             (:447)
            float2 b[7];
            ld14_v2(b, myT + m * S);                                   // column m of A
            axpy2(th0, th1, getp(phi0, m), getp(phi1, m), b);
        }
#pragma unroll
        for (int m = 0; m < 7; m++) {                                  // + BR B^T              (:473,:485)
            float2 b[7];
            ld14_v2(b, myP + m * PR + cL);                             // column m of B
            axpy2(th0, th1, br0[m], br1[m], b);
        }
        // gamma  (:416-443)
        float gm0, gm1;
        {
            float2 x[7];
            ld14_v4(x, vqm);
            const float2 s0 = dot14_acc(phi0, x, splat(0.f)), s1 = dot14_acc(phi1, x, splat(0.f));
            const float4 ra = reinterpret_cast<const float4*>(vrm)[0], rb = reinterpret_cast<const float4*>(vrm)[1];
            const float rv[7] = {ra.x, ra.y, ra.z, ra.w, rb.x, rb.y, rb.z};
            float d0 = 0.f, d1 = 0.f;
#pragma unroll
            for (int j = 0; j < 7; j++) { d0 = fmaf(br0[j], rv[j], d0); d1 = fmaf(br1[j], rv[j], d1); }
            gm0 = (ti0 - vc[r0]) + (d0 + (s0.x + s0.y));
            gm1 = (ti1 - vc[r1]) + (d1 + (s1.x + s1.y));
        }
        if (inner) {
#pragma unroll
            for (int q = 0; q < 7; q++) { R.Sd[0][q] = f2(-th0[q].x, -th0[q].y); R.Sd[1][q] = f2(-th1[q].x, -th1[q].y); }   // :501
            R.gamma[0] = -gm0; R.gamma[1] = -gm1;                                                    // :531
        } else {
#pragma unroll
            for (int q = 0; q < 7; q++) { R.Sd[0][q] = f2(-aQ0[q].x, -aQ0[q].y); R.Sd[1][q] = f2(-aQ1[q].x, -aQ1[q].y); }   // :249
            R.gamma[0] = -ti0; R.gamma[1] = -ti1;                                                    // :274
        }
        // phi rows replace A in this knot's tile (only this knot's lanes read A): S[k][0] = -phi (:491)
        __syncwarp();
        if (k != 0 && act) { st14_v2(myT + r0 * S, phi0); st14_v2(myT + r1 * S, phi1); }
    }
    gj14(th0, th1, t, act, piv);                                       // theta_k^-1            (:511-513)

    cluster_sync_all<CL>();                        // every knot (and every CTA's halo pull) is done with Qt^-1
    // Pd rows: -theta_k^-1 (:518-524), knot 0: -Qt_0 (:202-210)
    if (k == 0) { ld14_v2(th0, myT + r0 * S); ld14_v2(th1, myT + r1 * S); }
    if (act) {
        float2 n0[7], n1[7];
#pragma unroll
        for (int q = 0; q < 7; q++) { n0[q] = f2(-th0[q].x, -th0[q].y); n1[q] = f2(-th1[q].x, -th1[q].y); }
        st14_v4(myP + r0 * PR + cD, n0);
        st14_v4(myP + r1 * PR + cD, n1);
    }
    cluster_sync_all<CL>();
    if (CL > 1) {
        if (rank > 0) pull_pd<CF>(smem, 0, KPC, rank - 1);
        if (rank + 1 < (unsigned)CL) pull_floats<CF>(TA + (KPC + 1) * TK, TA + 1 * TK, TK, rank + 1);
        __syncthreads();
    }

    // ---- phase 2: Pinv left block = theta_k^-1 phi_k theta_{k-1}^-1 (stored sign, :97-115) ----
    {
        float2 t10[7], t11[7], pl0[7], pl1[7];
#pragma unroll
        for (int q = 0; q < 7; q++) { t10[q] = splat(0.f); t11[q] = splat(0.f); pl0[q] = splat(0.f); pl1[q] = splat(0.f); }
#pragma unroll
        for (int m = 0; m < 14; m++) {                                 // theta_k^-1 phi_k        (:100)
            float2 b[7];
            ld14_v2(b, myT + m * S);
            axpy2(t10, t11, getp(th0, m), getp(th1, m), b);
        }
        const float* Pm = PW + (sl - 1) * PK + cD;                     // -theta_{k-1}^-1 rows
#pragma unroll
        for (int m = 0; m < 14; m++) {                                 // ... theta_{k-1}^-1      (:102)
            float2 b[7];
            ld14_v4(b, Pm + m * PR);
            axpy2(pl0, pl1, -getp(t10, m), -getp(t11, m), b);
        }
        if (!inner) {
#pragma unroll
            for (int q = 0; q < 7; q++) { pl0[q] = splat(0.f); pl1[q] = splat(0.f); }
        }
        // own rows -> Pl columns; transposed -> Pr columns of the left neighbour (:118-136 by symmetry)
        if (act) {
            st14_v2(myP + r0 * PR + cL, pl0);
            st14_v2(myP + r1 * PR + cL, pl1);
            float* left = PW + (sl - 1) * PK + cR;
            if (CL > 1 && g == 0 && rank > 0) left = map_rank<CL>(PW + KPC * PK + cR, rank - 1);
#pragma unroll
            for (int j = 0; j < 14; j++) { left[j * PR + r0] = getp(pl0, j); left[j * PR + r1] = getp(pl1, j); }
        }
    }
    // last valid knot (and padding knots): no right neighbour writes their Pr columns
    if (!has_right && act) {
#pragma unroll
        for (int j = 0; j < 14; j++) { myP[j * PR + cR + r0] = 0.f; myP[j * PR + cR + r1] = 0.f; }
    }
    // ---- S left rows (own phi rows) and right rows (-phi_{k+1}^T, :537-557) ------------------
    {
        float2 a[7], b[7];
        ld14_v2(a, myT + r0 * S); ld14_v2(b, myT + r1 * S);
#pragma unroll
        for (int q = 0; q < 7; q++) {
            R.Sl[0][q] = inner ? f2(-a[q].x, -a[q].y) : splat(0.f);
            R.Sl[1][q] = inner ? f2(-b[q].x, -b[q].y) : splat(0.f);
        }
        const float* phin = TA + (sl + 1) * TK;
#pragma unroll
        for (int j = 0; j < 14; j++) {
            setp(R.Sr[0], j, has_right ? -phin[j * S + r0] : 0.f);
            setp(R.Sr[1], j, has_right ? -phin[j * S + r1] : 0.f);
        }
    }
    if (!valid || !act) {
#pragma unroll
        for (int q = 0; q < 7; q++) {
            R.Sl[0][q] = R.Sl[1][q] = R.Sd[0][q] = R.Sd[1][q] = R.Sr[0][q] = R.Sr[1][q] = splat(0.f);
        }
        R.gamma[0] = R.gamma[1] = 0.f;
    }
    cluster_sync_all<CL>();                        // PROW complete (incl. rows scattered by neighbours)
    if (!valid && act) {                           // padding knots: zero Pinv rows (their Pd holds 1/rho junk)
#pragma unroll
        for (int q = 0; q < 11; q++) {
            reinterpret_cast<float4*>(myP + r0 * PR)[q] = make_float4(0.f, 0.f, 0.f, 0.f);
            reinterpret_cast<float4*>(myP + r1 * PR)[q] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
}

// ---------------------------------------------------------------------------------------
// Cluster-wide sum with a bit-identical result in every thread of every CTA: warp shuffle,
// one (D)SMEM store per (warp, destination CTA), ONE barrier, then every thread adds the same
// NRED partials in the same order.
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
    const float4* r4 = reinterpret_cast<const float4*>(red + buf * CF::NREDP);
    float acc = 0.f;
#pragma unroll
    for (int q = 0; q < CF::NREDP / 4; q++) {
        const float4 v = r4[q];
        acc += v.x;
        if (4 * q + 1 < NRED) acc += v.y;
        if (4 * q + 2 < NRED) acc += v.z;
        if (4 * q + 3 < NRED) acc += v.w;
    }
    return acc;
}

// ---------------------------------------------------------------------------------------
// pcg_rows: preconditioned CG on the stored system S lambda = gamma.  Semantics (warm start,
// |r.Pinv r| < tol absolute exit, iteration count, max-iter flag) follow SURVEY 0.2 /
// oracle_pcg.  Per iteration: 2 reductions (one barrier each: __syncthreads for CL == 1, a
// cluster barrier otherwise) + 2 CTA barriers; the boundary rows of S p and Pinv r ride
// along with the partial sums, so neighbouring CTAs update their halo copies of r and p
// locally and need no further exchange.
// ---------------------------------------------------------------------------------------
template <class CF>
__device__ __forceinline__ void pcg_rows(float* smem, const Rows& R, float* lam_g, int N,
                                         unsigned max_iter, float exit_tol, unsigned* d_iters,
                                         uint8_t* d_exit) {
    constexpr int S = 14, KPC = CF::KPC, CL = CF::CL, PR = CF::PR, PK = CF::PK, VK = CF::VK;
    const int tid = threadIdx.x;
    const int g = tid >> 3, t = tid & 7;
    const unsigned rank = cluster_rank<CL>();
    const int k0 = rank * KPC, k = k0 + g;
    const bool act = t < 7, valid = k < N;
    const int r0 = act ? t : 0, r1 = r0 + 7;
    const bool first_g = (g == 0), last_g = (g == KPC - 1);
    const bool has_l = rank > 0, has_r = rank + 1 < (unsigned)CL;
    const int sl = g + 1;

    float* pv = smem + CF::oPv;
    float* rv = smem + CF::oRv;
    float* hal = smem + CF::oHal;   // [phase][side][16]; side 0: pushed by my LEFT neighbour CTA
    float* red = smem + CF::oRed;
    const float* myP0 = smem + CF::oP + sl * PK + r0 * PR;
    const float* myP1 = smem + CF::oP + sl * PK + r1 * PR;

    // y(rows r0, r1) = [L D R] . [x_{k-1}; x_k; x_{k+1}], S rows from registers
    auto s_mv = [&](const float* vec, float& y0, float& y1) {
        float2 xl[7], xd[7], xr[7];
        ld14_v4(xl, vec + (sl - 1) * VK); ld14_v4(xd, vec + sl * VK); ld14_v4(xr, vec + (sl + 1) * VK);
        float2 a0 = dot14_acc(R.Sl[0], xl, splat(0.f)), b0 = dot14_acc(R.Sd[0], xd, splat(0.f)), c0 = dot14_acc(R.Sr[0], xr, splat(0.f));
        float2 a1 = dot14_acc(R.Sl[1], xl, splat(0.f)), b1 = dot14_acc(R.Sd[1], xd, splat(0.f)), c1 = dot14_acc(R.Sr[1], xr, splat(0.f));
        y0 = ((a0.x + a0.y) + (b0.x + b0.y)) + (c0.x + c0.y);
        y1 = ((a1.x + a1.y) + (b1.x + b1.y)) + (c1.x + c1.y);
    };
    // same with the Pinv rows from shared memory (row = [Pd | Pl | Pr | pad])
    auto p_mv = [&](const float* vec, float& y0, float& y1) {
        float2 xl[7], xd[7], xr[7];
        ld14_v4(xl, vec + (sl - 1) * VK); ld14_v4(xd, vec + sl * VK); ld14_v4(xr, vec + (sl + 1) * VK);
        const float4* q0 = reinterpret_cast<const float4*>(myP0);
        const float4* q1 = reinterpret_cast<const float4*>(myP1);
        float2 acc0[3] = {splat(0.f), splat(0.f), splat(0.f)}, acc1[3] = {splat(0.f), splat(0.f), splat(0.f)};
#pragma unroll
        for (int q = 0; q < 11; q++) {
            const float4 u = q0[q], v = q1[q];
            // floats 4q .. 4q+3 of the row: pair indices 2q and 2q+1 of the 21 pairs (Pd 0-6, Pl 7-13, Pr 14-20)
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const int pi = 2 * q + h;
                if (pi < 21) {
                    const int part = pi / 7, w = pi % 7;
                    const float2 x = (part == 0) ? xd[w] : (part == 1) ? xl[w] : xr[w];
                    acc0[part] = fma2(h ? f2(u.z, u.w) : f2(u.x, u.y), x, acc0[part]);
                    acc1[part] = fma2(h ? f2(v.z, v.w) : f2(v.x, v.y), x, acc1[part]);
                }
            }
        }
        y0 = ((acc0[1].x + acc0[1].y) + (acc0[0].x + acc0[0].y)) + (acc0[2].x + acc0[2].y);
        y1 = ((acc1[1].x + acc1[1].y) + (acc1[0].x + acc1[0].y)) + (acc1[2].x + acc1[2].y);
    };

    float* my_p0 = pv + sl * VK + r0; float* my_p1 = pv + sl * VK + r1;
    float* my_r0 = rv + sl * VK + r0; float* my_r1 = rv + sl * VK + r1;
    float* hl_p = pv;                        // left halo slot
    float* hr_p = pv + (KPC + 1) * VK;       // right halo slot
    float* hl_r = rv;
    float* hr_r = rv + (KPC + 1) * VK;

    // ---- r = gamma - S lambda0 --------------------------------------------------------------
    float lam0 = 0.f, lam1 = 0.f;
    if (valid && act) { lam0 = lam_g[(size_t)k * S + r0]; lam1 = lam_g[(size_t)k * S + r1]; }
    if (act) { *my_p0 = lam0; *my_p1 = lam1; }
    if (t == 7) { pv[sl * VK + 14] = 0.f; pv[sl * VK + 15] = 0.f; rv[sl * VK + 14] = 0.f; rv[sl * VK + 15] = 0.f; }
    if (first_g) {
        const bool ok = (k0 >= 1 && k0 - 1 < N && act);
        if (act) { hl_p[r0] = ok ? lam_g[(size_t)(k0 - 1) * S + r0] : 0.f; hl_p[r1] = ok ? lam_g[(size_t)(k0 - 1) * S + r1] : 0.f; }
        else { hl_p[14] = hl_p[15] = 0.f; hl_r[14] = hl_r[15] = 0.f; }
    }
    if (last_g) {
        const bool ok = (k0 + KPC < N && act);
        if (act) { hr_p[r0] = ok ? lam_g[(size_t)(k0 + KPC) * S + r0] : 0.f; hr_p[r1] = ok ? lam_g[(size_t)(k0 + KPC) * S + r1] : 0.f; }
        else { hr_p[14] = hr_p[15] = 0.f; hr_r[14] = hr_r[15] = 0.f; }
    }
    __syncthreads();
    float y0, y1;
    s_mv(pv, y0, y1);
    float rr0 = R.gamma[0] - y0, rr1 = R.gamma[1] - y1;
    if (act) {
        *my_r0 = rr0; *my_r1 = rr1;
        if (first_g) { if (has_l) { float* d = map_rank<CL>(hr_r, rank - 1); d[r0] = rr0; d[r1] = rr1; } else { hl_r[r0] = 0.f; hl_r[r1] = 0.f; } }
        if (last_g)  { if (has_r) { float* d = map_rank<CL>(hl_r, rank + 1); d[r0] = rr0; d[r1] = rr1; } else { hr_r[r0] = 0.f; hr_r[r1] = 0.f; } }
    }
    cluster_sync_all<CL>();

    // ---- p = rt = Pinv r ; eta = r.rt ----------------------------------------------------------
    float rt0, rt1;
    p_mv(rv, rt0, rt1);
    float p0 = rt0, p1 = rt1;
    if (act) {
        *my_p0 = p0; *my_p1 = p1;       // pv was last read before the barrier above
        if (first_g) { if (has_l) { float* d = map_rank<CL>(hr_p, rank - 1); d[r0] = p0; d[r1] = p1; } else { hl_p[r0] = 0.f; hl_p[r1] = 0.f; } }
        if (last_g)  { if (has_r) { float* d = map_rank<CL>(hl_p, rank + 1); d[r0] = p0; d[r1] = p1; } else { hr_p[r0] = 0.f; hr_p[r1] = 0.f; } }
    }
    float eta = cluster_sum<CF>(red, 1, act ? fmaf(rr0, rt0, rr1 * rt1) : 0.f, rank);

    unsigned iter = 0;
    bool max_iter_exit = true;
    for (; iter < max_iter; iter++) {
        // ---- ups = S p ; v = p.ups -----------------------------------------------------------
        float u0, u1;
        s_mv(pv, u0, u1);
        if (CL > 1 && act) {
            if (first_g && has_l) { float* d = map_rank<CL>(hal + (0 * 2 + 1) * VK, rank - 1); d[r0] = u0; d[r1] = u1; }
            if (last_g && has_r)  { float* d = map_rank<CL>(hal + (0 * 2 + 0) * VK, rank + 1); d[r0] = u0; d[r1] = u1; }
        }
        const float v = cluster_sum<CF>(red, 0, act ? fmaf(p0, u0, p1 * u1) : 0.f, rank);
        if (v == 0.f) { max_iter_exit = false; break; }
        const float alpha = eta / v;
        lam0 = fmaf(alpha, p0, lam0); lam1 = fmaf(alpha, p1, lam1);
        rr0 = fmaf(-alpha, u0, rr0); rr1 = fmaf(-alpha, u1, rr1);
        if (act) {
            *my_r0 = rr0; *my_r1 = rr1;
            if (CL > 1) {
                if (first_g && has_l) { hl_r[r0] = fmaf(-alpha, hal[(0 * 2 + 0) * VK + r0], hl_r[r0]); hl_r[r1] = fmaf(-alpha, hal[(0 * 2 + 0) * VK + r1], hl_r[r1]); }
                if (last_g && has_r)  { hr_r[r0] = fmaf(-alpha, hal[(0 * 2 + 1) * VK + r0], hr_r[r0]); hr_r[r1] = fmaf(-alpha, hal[(0 * 2 + 1) * VK + r1], hr_r[r1]); }
            }
        }
        __syncthreads();
        // ---- rt = Pinv r ; eta_new = r.rt ----------------------------------------------------
        p_mv(rv, rt0, rt1);
        if (CL > 1 && act) {
            if (first_g && has_l) { float* d = map_rank<CL>(hal + (1 * 2 + 1) * VK, rank - 1); d[r0] = rt0; d[r1] = rt1; }
            if (last_g && has_r)  { float* d = map_rank<CL>(hal + (1 * 2 + 0) * VK, rank + 1); d[r0] = rt0; d[r1] = rt1; }
        }
        const float eta_new = cluster_sum<CF>(red, 1, act ? fmaf(rr0, rt0, rr1 * rt1) : 0.f, rank);
        if (fabsf(eta_new) < exit_tol) { iter++; max_iter_exit = false; break; }
        const float beta = eta_new / eta;
        eta = eta_new;
        p0 = fmaf(beta, p0, rt0); p1 = fmaf(beta, p1, rt1);
        if (act) {
            *my_p0 = p0; *my_p1 = p1;
            if (CL > 1) {
                if (first_g && has_l) { hl_p[r0] = fmaf(beta, hl_p[r0], hal[(1 * 2 + 0) * VK + r0]); hl_p[r1] = fmaf(beta, hl_p[r1], hal[(1 * 2 + 0) * VK + r1]); }
                if (last_g && has_r)  { hr_p[r0] = fmaf(beta, hr_p[r0], hal[(1 * 2 + 1) * VK + r0]); hr_p[r1] = fmaf(beta, hr_p[r1], hal[(1 * 2 + 1) * VK + r1]); }
            }
        }
        __syncthreads();
    }
    if (valid && act) { lam_g[(size_t)k * S + r0] = lam0; lam_g[(size_t)k * S + r1] = lam1; }
    if (rank == 0 && tid == 0) { *d_iters = iter; *d_exit = max_iter_exit ? 1 : 0; }
    // no CTA may exit while a peer can still address its shared memory
    if (CL > 1) cluster_sync_all<CL>();
}

}  // namespace mpcb200
