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
// Mapping ("row owner"): LPK lanes per knot (4 or 8); lane t owns rows t, t+LPK, t+2*LPK, ...
// (< 14) of every 14 x 14 block of its knot: RPT = 4 or 2 row slots per lane, 14 of 16 slots
// used.  32/LPK knots share a warp, so everything inside a knot synchronises with
// __syncwarp().  One CTA holds up to 64 knots; a whole 64-knot instance lives in ONE SM:
//   * S block rows (3 x 14 floats per row) stay in registers for the whole solve;
//   * Pinv block rows live in shared memory as PROW[knot][row][44] = [Pd | Pl | Pr | pad2]
//     (row stride 44 floats, knot stride 624: 16 B aligned and bank-conflict free);
//   * all pair-wise arithmetic is FFMA2 (fma.rn.f32x2, sm_100): half the FMA issue slots.
// The PCG iteration is bound by the shared-memory pipe (128 B/clk/SM): every x value a lane
// needs is one register delivery, every Pinv coefficient read from shared memory is 4 B per
// FMA.  LPK = 4 (4 rows per lane) halves the x deliveries per FMA relative to LPK = 8.
// The Schur phase uses the same shared memory as scratch: the Pd columns of PROW first hold
// Qt^-1 rows, then -theta^-1 rows; the (not yet written) Pl / Pr columns hold B, R^-1 and
// the knot's vectors; a second tile per knot holds A, then phi.
// Longer horizons use a thread-block cluster of CL CTAs with DSMEM halos.
// Everything is templated on the scalar type through Cfg<KPC, CL, LPK, TM, T>: T = float (the reference's
// default linsys_t; pairs are float2 + FFMA2, Pinv rows may live in tensor memory) or T = double
// (USE_DOUBLES 1; pairs are double2 + two DFMAs, Pinv rows stay in shared memory, LPK may be 16).
#pragma once
#include "common.cuh"

namespace mpcb200 {

#ifndef MPCB200_PINV_TMEM
#define MPCB200_PINV_TMEM 1   // 0: Pinv rows stay in shared memory during the PCG phase (A/B builds)
#endif
#ifndef MPCB200_PCG_VARIANT
// iteration of pcg_rows: 0 = the two-reduction PCG as the reference states it (the default, every tiling);
//   1 = single-reduction PCG (Chronopoulos-Gear; fails the +-1 iteration criterion on two real-IIWA solves);
//   2 = S p carried by a recurrence so that the r.rt reduction hides behind the matvec S rt (single-CTA tilings; cluster and
//       knot-split tilings run 0).  4 % faster on the bench workload and within the operating-point criteria, but measured
//       5-20x farther from the exact-arithmetic iterate than variant 0 at tight tolerances (tools/variant_drift.py,
//       profiles/r2i_variant_drift_table.txt): the recurrence loses what the matvec S p would recompute.
// 1 and 2 are A/B build options only; parity comes first.
#define MPCB200_PCG_VARIANT 0
#endif
constexpr int pow2_at_least(int x) { int p = 32; while (p < x) p *= 2; return p; }

// scalar traits: T2 = the 16 B (double) / 8 B (float) pair type all pair-wise arithmetic runs on
template <class T> struct Num;
template <> struct Num<float>  { using T2 = float2; };
template <> struct Num<double> { using T2 = double2; };

// TM_ = 1: the PCG phase keeps the Pinv rows in tensor memory instead of shared memory (fp32 only)
// T_: float (settings.cuh USE_DOUBLES 0, the reference default) or double (USE_DOUBLES 1)
template <int KPC_, int CL_, int LPK_, int TM_ = MPCB200_PINV_TMEM, class T_ = float>
struct Cfg {
    using T = T_;
    using T2 = typename Num<T_>::T2;
    static constexpr bool F32 = sizeof(T_) == 4;
    static constexpr bool TM = TM_ != 0 && F32;
    static constexpr int S = 14, C = 7;
    static constexpr int KPC = KPC_, CL = CL_;
    static constexpr int LPK = LPK_;            // lanes per knot
    static constexpr int RPT = (S + LPK - 1) / LPK;   // row slots per lane (14 x 14 blocks)
    static constexpr int RPT7 = (C + LPK - 1) / LPK;  // row slots per lane (7 x 7 block)
    static constexpr int NT = KPC * LPK;        // threads per CTA
    static constexpr int NW = NT / 32;          // warps per CTA
    static constexpr int SLOTS = KPC + 2;       // one halo knot on both sides
    static constexpr int PR = 44;               // floats per PROW row
    static constexpr int PK = S * PR + 8;       // floats per PROW knot (624 = 16 mod 32 banks)
    static constexpr int TK = S * S;            // floats per A / phi tile (196)
    static constexpr int VK = 20;               // floats per vector slot (14 used; 80 B stride: the 4 or 8
                                                // knots a warp reads land in distinct banks)
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
    static constexpr int oMb = oRed + 2 * NREDP;        // 2 mbarriers (8 B each), cluster kernels
    static constexpr int oTm = oMb + 4;                 // TMEM base address of this CTA (+ pad)
    static constexpr int FLOATS = oTm + 4;
    // tensor-memory copy of Sr | Pl | Pr: per lane 7 chunks of 2*RPT columns (Sr pair w of my RPT rows) and 7 chunks of
    // 4*RPT (Pl pair w | Pr pair w)
    static constexpr int TPT = 42 * RPT;                // columns per thread
    static constexpr int TCOLS = pow2_at_least(((NW + 3) / 4) * TPT);   // allocation, power of two <= 512
    static_assert(!TM || TCOLS <= 512, "tensor memory has 512 columns");
    static constexpr size_t SMEM_BYTES = sizeof(T) * FLOATS;
    static constexpr int MIN_CTAS = !F32 ? (SMEM_BYTES <= 75 * 1024 && NT <= 64 ? 3 : SMEM_BYTES <= 113 * 1024 && NT <= 128 ? 2 : 1) : (LPK == 4) ? (256 / NT > 0 ? 256 / NT : 1) : (512 / NT > 0 ? 512 / NT : 1);
    static_assert(LPK == 4 || LPK == 8 || LPK == 16, "4, 8 or 16 lanes per knot");
    static_assert(NT % 32 == 0 && NT <= 1024, "whole warps, <= 1024 threads");
    static_assert(SMEM_BYTES <= 232448, "exceeds the 227 KB per-CTA shared memory of sm_100");
};

// Per-instance views of the reference buffers (include/pcg/sqp.cuh:54-58).
template <class T>
struct KktPtrs {
    T* G;         // in: [Q_k | R_k]; out: [Qt_k^-1 | Rt_k^-1]
    const T* C;   // [Abar_k | Bbar_k]
    const T* g;   // [q_k | r_k]
    const T* c;   // constraint residuals
};

// Register-resident part of a knot's block rows: the lane's RPT rows of the stored (negated)
// Schur complement, as pairs.
template <class T, int RPT>
struct Rows {
    typename Num<T>::T2 Sl[RPT][7], Sd[RPT][7], Sr[RPT][7];
    T gamma[RPT];
};
template <class CF> using RowsOf = Rows<typename CF::T, CF::RPT>;

// ---- small pair helpers (float2: FFMA2; double2: two DFMAs) ---------------------------------
__device__ __forceinline__ float2 f2(float a, float b) { return make_float2(a, b); }
__device__ __forceinline__ float2 splat(float a) { return make_float2(a, a); }
__device__ __forceinline__ float2 neg2(float2 a) { return make_float2(-a.x, -a.y); }
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ double2 f2(double a, double b) { return make_double2(a, b); }
__device__ __forceinline__ double2 splat(double a) { return make_double2(a, a); }
__device__ __forceinline__ double2 neg2(double2 a) { return make_double2(-a.x, -a.y); }
__device__ __forceinline__ double2 fma2(double2 a, double2 b, double2 c) { return make_double2(fma(a.x, b.x, c.x), fma(a.y, b.y, c.y)); }
__device__ __forceinline__ float fm(float a, float b, float c) { return fmaf(a, b, c); }
__device__ __forceinline__ double fm(double a, double b, double c) { return fma(a, b, c); }
#ifndef MPCB200_FAST_DIV
#define MPCB200_FAST_DIV 0    // 1 (experiment): alpha, beta by the approximate division (2 ulp) instead of the IEEE one
#endif
__device__ __forceinline__ float divT(float a, float b) { return MPCB200_FAST_DIV ? __fdividef(a, b) : a / b; }
__device__ __forceinline__ double divT(double a, double b) { return a / b; }
__device__ __forceinline__ float absT(float a) { return fabsf(a); }
__device__ __forceinline__ double absT(double a) { return fabs(a); }
__device__ __forceinline__ float getp(const float2 (&a)[7], int i) { return (i & 1) ? a[i >> 1].y : a[i >> 1].x; }
__device__ __forceinline__ void setp(float2 (&a)[7], int i, float v) { if (i & 1) a[i >> 1].y = v; else a[i >> 1].x = v; }
__device__ __forceinline__ double getp(const double2 (&a)[7], int i) { return (i & 1) ? a[i >> 1].y : a[i >> 1].x; }
__device__ __forceinline__ void setp(double2 (&a)[7], int i, double v) { if (i & 1) a[i >> 1].y = v; else a[i >> 1].x = v; }

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
// doubles: every 14-element run starts at an even element offset, i.e. 16 B aligned -> seven 16 B accesses
__device__ __forceinline__ void ld14_v2(double2 (&r)[7], const double* p) {
    const double2* s = reinterpret_cast<const double2*>(p);
#pragma unroll
    for (int q = 0; q < 7; q++) r[q] = s[q];
}
__device__ __forceinline__ void ld14_v4(double2 (&r)[7], const double* p) { ld14_v2(r, p); }
__device__ __forceinline__ void st14_v2(double* p, const double2 (&r)[7]) {
    double2* d = reinterpret_cast<double2*>(p);
#pragma unroll
    for (int q = 0; q < 7; q++) d[q] = r[q];
}
__device__ __forceinline__ void st14_v4(double* p, const double2 (&r)[7]) { st14_v2(p, r); }
// 8 elements (a 7-row padded to 8) to / from a 16 B aligned address
__device__ __forceinline__ void st8(float* p, const float (&a)[8]) {
    reinterpret_cast<float4*>(p)[0] = make_float4(a[0], a[1], a[2], a[3]);
    reinterpret_cast<float4*>(p)[1] = make_float4(a[4], a[5], a[6], a[7]);
}
__device__ __forceinline__ void ld8(float (&a)[8], const float* p) {
    const float4 u = reinterpret_cast<const float4*>(p)[0], v = reinterpret_cast<const float4*>(p)[1];
    a[0] = u.x; a[1] = u.y; a[2] = u.z; a[3] = u.w; a[4] = v.x; a[5] = v.y; a[6] = v.z; a[7] = v.w;
}
__device__ __forceinline__ void st8(double* p, const double (&a)[8]) {
#pragma unroll
    for (int q = 0; q < 4; q++) reinterpret_cast<double2*>(p)[q] = make_double2(a[2 * q], a[2 * q + 1]);
}
__device__ __forceinline__ void ld8(double (&a)[8], const double* p) {
#pragma unroll
    for (int q = 0; q < 4; q++) { const double2 v = reinterpret_cast<const double2*>(p)[q]; a[2 * q] = v.x; a[2 * q + 1] = v.y; }
}

// dot of a 14-row (pairs) with 14 values: two interleaved partial sums (even / odd)
template <class T2>
__device__ __forceinline__ T2 dot14_acc(const T2 (&a)[7], const T2 (&x)[7], T2 acc) {
#pragma unroll
    for (int q = 0; q < 7; q++) acc = fma2(a[q], x[q], acc);
    return acc;
}
__device__ __forceinline__ float hsum(float2 a) { return a.x + a.y; }
__device__ __forceinline__ double hsum(double2 a) { return a.x + a.y; }

// c[j] += a[j] * b for the lane's RPT rows (one broadcast row b shared by all of them)
template <int RPT, class T, class T2>
__device__ __forceinline__ void axpy_rows(T2 (&c)[RPT][7], const T (&a)[RPT], const T2 (&b)[7]) {
#pragma unroll
    for (int j = 0; j < RPT; j++) {
        const T2 s = splat(a[j]);
#pragma unroll
        for (int q = 0; q < 7; q++) c[j][q] = fma2(s, b[q], c[j][q]);
    }
}

// Row bookkeeping of a lane: slot j holds row t + LPK*j (rows >= D do not exist; such slots
// shadow lane 0's row of that slot for addressing (same address: a broadcast, no bank
// conflict), never store, and carry zeros into every sum).
template <int LPK, int RPT, int D>
struct RowMap {
    int row[RPT];
    bool ok[RPT];
    __device__ __forceinline__ RowMap(int t) {
#pragma unroll
        for (int j = 0; j < RPT; j++) { const int r = t + LPK * j; ok[j] = r < D; row[j] = ok[j] ? r : LPK * j; }
    }
};

// In-place Gauss-Jordan inverse WITHOUT pivoting of a 14 x 14 matrix held as rows (lane t
// owns rows t + LPK*j).  Same elimination sequence as the reference's invertMatrix on [A | I]
// (include/utils/matrix.cuh:120-238): the in-place form keeps the evolving inverse column in
// the slot of the eliminated one.  `piv`: 16 floats of shared memory private to this knot.
template <int LPK, int RPT, class T2, class T>
__device__ __forceinline__ void gj14(T2 (&a)[RPT][7], int t, T* piv) {
#pragma unroll
    for (int p = 0; p < 14; p++) {
        const int owner = p % LPK, slot = p / LPK;
        if (t == owner) st14_v4(piv, a[slot]);
        __syncwarp();
        T2 rp[7];
        ld14_v4(rp, piv);
        __syncwarp();
        const T pinv = T(1) / getp(rp, p);
#pragma unroll
        for (int j = 0; j < RPT; j++) {
            const T f = getp(a[j], p) * pinv;
            if (j == slot && t == owner) {
#pragma unroll
                for (int q = 0; q < 7; q++) a[j][q] = f2(rp[q].x * pinv, rp[q].y * pinv);
                setp(a[j], p, pinv);
            } else {
                const T2 nf = splat(-f);
#pragma unroll
                for (int q = 0; q < 7; q++) a[j][q] = fma2(nf, rp[q], a[j][q]);
                setp(a[j], p, -f);
            }
        }
    }
}

// 7 x 7 Gauss-Jordan, rows t + LPK*j (8 values per row, last one padding).  piv: 8 values.
template <int LPK, int RPT7, class T>
__device__ __forceinline__ void gj7(T (&a)[RPT7][8], int t, T* piv) {
#pragma unroll
    for (int p = 0; p < 7; p++) {
        const int owner = p % LPK, slot = p / LPK;
        if (t == owner) {
            const T row[8] = {a[slot][0], a[slot][1], a[slot][2], a[slot][3], a[slot][4], a[slot][5], a[slot][6], T(0)};
            st8(piv, row);
        }
        __syncwarp();
        T rp[8];
        ld8(rp, piv);
        __syncwarp();
        const T pinv = T(1) / rp[p];
#pragma unroll
        for (int j = 0; j < RPT7; j++) {
            const T f = a[j][p] * pinv;
            if (j == slot && t == owner) {
#pragma unroll
                for (int c = 0; c < 7; c++) a[j][c] = rp[c] * pinv;
                a[j][p] = pinv;
            } else {
#pragma unroll
                for (int c = 0; c < 7; c++)
                    if (c != p) a[j][c] = fm(-f, rp[c], a[j][c]);
                a[j][p] = -f;
            }
        }
    }
}

// copy one neighbour-CTA region into local shared memory (cluster kernels only)
template <class CF>
__device__ __forceinline__ void pull_floats(typename CF::T* dst, typename CF::T* src_local_addr, int n, unsigned src_rank) {
    using T2 = typename CF::T2;
    const T2* src = reinterpret_cast<const T2*>(map_rank<CF::CL>(src_local_addr, src_rank));
    T2* d = reinterpret_cast<T2*>(dst);
    for (int q = threadIdx.x; q < n / 2; q += CF::NT) d[q] = src[q];
}
// Pd columns (14 x 14, row stride PR) of one PROW knot from a neighbour CTA
template <class CF>
__device__ __forceinline__ void pull_pd(typename CF::T* smem, int dst_slot, int src_slot, unsigned src_rank) {
    using T = typename CF::T;
    using T2 = typename CF::T2;
    const T* src = map_rank<CF::CL>(smem + CF::oP + src_slot * CF::PK, src_rank);
    T* dst = smem + CF::oP + dst_slot * CF::PK;
    for (int q = threadIdx.x; q < 14 * 7; q += CF::NT) {
        const int row = q / 7, c2 = q - row * 7;
        reinterpret_cast<T2*>(dst + row * CF::PR)[c2] = reinterpret_cast<const T2*>(src + row * CF::PR)[c2];
    }
}

// ---------------------------------------------------------------------------------------
// schur_rows: everything form_schur_system produces, for the knots owned by this CTA.
// Leaves S rows and gamma in registers and the Pinv rows in PROW; writes G^-1 in place.
// ---------------------------------------------------------------------------------------
template <class CF>
__device__ __forceinline__ void schur_rows(typename CF::T* smem, const KktPtrs<typename CF::T>& P, int N,
                                           typename CF::T rho, RowsOf<CF>& R) {
    using T = typename CF::T;
    using T2 = typename CF::T2;
    constexpr int S = 14, C = 7, KPC = CF::KPC, CL = CF::CL, PR = CF::PR, PK = CF::PK, TK = CF::TK;
    constexpr int LPK = CF::LPK, RPT = CF::RPT, RPT7 = CF::RPT7;
    constexpr int SS = S * S, CC = C * C, SC = S * C;
    constexpr int GSET = SS + CC, CSET = SS + SC, gSET = S + C;
    constexpr int cD = CF::cD, cL = CF::cL, cR = CF::cR;
    const int tid = threadIdx.x;
    const int g = tid / LPK, t = tid % LPK;
    const unsigned rank = cluster_rank<CL>();
    const int k0 = rank * KPC;
    const int k = k0 + g;
    const RowMap<LPK, RPT, 14> rm(t);
    const RowMap<LPK, RPT7, 7> rm7(t);
    const bool valid = k < N;
    const bool inner = valid && k >= 1;
    const bool has_right = valid && (k + 1 < N);
    const int sl = g + 1;

    T* PW = smem + CF::oP;                    // PROW
    T* TA = smem + CF::oT;
    T* myP = PW + sl * PK;
    T* myT = TA + sl * TK;
    // scratch living in the not-yet-written Pl / Pr columns of this knot's PROW rows
    T* vq = myP + 0 * PR + cR;                // q_k
    T* vqm = myP + 1 * PR + cR;               // q_{k-1}
    T* vc = myP + 2 * PR + cR;                // c_k
    T* vrm = myP + 3 * PR + cR;               // r_{k-1} (7, padded to 8)
    T* piv = myP + 4 * PR + cR;               // pivot row broadcast (16 B aligned, + 2 pad)
    // B_{k-1} column m: myP[m*PR + cL .. +13] (m < 7); Rt^-1 row m: myP[(7+m)*PR + cL .. +7]

    // ---- stage [Abar_{k-1} | Bbar_{k-1}] (one contiguous run per knot), coalesced ----------
    // (T2 granularity: CSET, SS and S are even, so a pair never straddles a column; fixed
    // trip count + unroll keeps 8 loads in flight per lane instead of one)
    {
        constexpr int TOT2 = KPC * CSET / 2, ITERS = (TOT2 + CF::NT - 1) / CF::NT;
        const T2* src = reinterpret_cast<const T2*>(P.C + ((ptrdiff_t)k0 - 1) * CSET);
        // loads first (branch-free, UN in flight per lane), stores after: with the store's branches
        // in between the compiler keeps ONE load in flight and the CTA waits out ITERS HBM round trips
        // latency tilings (<= 20 T2 per lane): the whole tile in flight at once; the 64-knot
        // throughput tiling measured best with 8 (1.23M vs 1.21M solves/s with all 37)
        constexpr int UN = ITERS <= 20 ? ITERS : 8;
        for (int it0 = 0; it0 < ITERS; it0 += UN) {
            T2 v[UN];
#pragma unroll
            for (int u = 0; u < UN; u++) {
                const int i2 = tid + (it0 + u) * CF::NT;
                const int kk = k0 + (2 * i2) / CSET;
                const bool ld = (it0 + u < ITERS) && i2 < TOT2 && kk >= 1 && kk < N;
                v[u] = f2(T(0), T(0));
                if (ld) v[u] = __ldg(src + i2);
            }
#pragma unroll
            for (int u = 0; u < UN; u++) {
                const int i2 = tid + (it0 + u) * CF::NT;
                if (it0 + u < ITERS && i2 < TOT2) {
                    const int idx = 2 * i2;
                    const int gg = idx / CSET, e = idx - gg * CSET;
                    if (e < SS) *reinterpret_cast<T2*>(TA + (gg + 1) * TK + e) = v[u];     // column-major A
                    else {
                        const int m = (e - SS) / S, i = (e - SS) - m * S;
                        *reinterpret_cast<T2*>(PW + (gg + 1) * PK + m * PR + cL + i) = v[u];
                    }
                }
            }
        }
    }
    // halo knots: keep every tile that an edge knot may read finite
    for (int q = tid; q < 2 * PK; q += CF::NT) PW[(q < PK) ? q : (KPC + 1) * PK + (q - PK)] = T(0);
    for (int q = tid; q < 2 * TK; q += CF::NT) TA[(q < TK) ? q : (KPC + 1) * TK + (q - TK)] = T(0);

    // ---- loads: Qt_k rows, Rt_{k-1} rows, vectors ------------------------------------------
    T2 aQ[RPT][7];
#pragma unroll
    for (int j = 0; j < RPT; j++) {
#pragma unroll
        for (int c = 0; c < 14; c++) {
            T v = T(0);
            if (valid && rm.ok[j]) v = P.G[(size_t)k * GSET + c * S + rm.row[j]];
            setp(aQ[j], c, v + ((c == t + LPK * j) ? rho : T(0)));       // linsys_setup.cuh:180,329-330
        }
    }
    T aR[RPT7][8];
#pragma unroll
    for (int j = 0; j < RPT7; j++) {
#pragma unroll
        for (int c = 0; c < 7; c++) {
            T v = T(0);
            if (inner && rm7.ok[j]) v = P.G[(size_t)(k - 1) * GSET + SS + c * C + rm7.row[j]];
            aR[j][c] = v + ((c == t + LPK * j) ? rho : T(0));            // :331
        }
        aR[j][7] = T(0);
    }
    // the vector entries are fetched BEFORE the barrier (their round trip overlaps the block loads above: fused 4096 x 64
    // 3.053 -> 3.024 ms) and parked in shared memory after it (the vectors overwrite chunks the staging loop zeroes).
    // Issuing the G row loads ahead of the staging loop as well was slower (Schur phase 0.875 -> 0.896 ms).
    T gq[RPT], gqm[RPT], gc[RPT], grm[RPT7];
#pragma unroll
    for (int j = 0; j < RPT; j++) {
        const int r = rm.row[j];
        gq[j] = (rm.ok[j] && valid) ? __ldg(P.g + (size_t)k * gSET + r) : T(0);
        gqm[j] = (rm.ok[j] && inner) ? __ldg(P.g + (size_t)(k - 1) * gSET + r) : T(0);
        gc[j] = (rm.ok[j] && inner) ? __ldg(P.c + (size_t)k * S + r) : T(0);
    }
#pragma unroll
    for (int j = 0; j < RPT7; j++) grm[j] = (rm7.ok[j] && inner) ? __ldg(P.g + (size_t)(k - 1) * gSET + S + rm7.row[j]) : T(0);
    __syncthreads();                               // staging done
#pragma unroll
    for (int j = 0; j < RPT; j++) {
        if (rm.ok[j]) {
            const int r = rm.row[j];
            vq[r] = gq[j];
            vqm[r] = gqm[j];
            vc[r] = gc[j];
        }
    }
#pragma unroll
    for (int j = 0; j < RPT7; j++)
        if (rm7.ok[j]) vrm[rm7.row[j]] = grm[j];
    if (t == 0) vrm[7] = T(0);
    // knot 0: Pinv[0][1] = -Qt_0, stored BEFORE inversion (linsys_setup.cuh:202-210); its A
    // tile is unused, park the rows there
    if (k == 0) {
#pragma unroll
        for (int j = 0; j < RPT; j++) if (rm.ok[j]) st14_v2(myT + rm.row[j] * S, aQ[j]);
    }

    // ---- phase 1a: inverses ---------------------------------------------------------------
    gj14<LPK, RPT>(aQ, t, piv);                                        // Qt_k^-1
    gj7<LPK, RPT7>(aR, t, piv);                                        // Rt_{k-1}^-1   (the two loops merged into one pivot loop,
                                                                       //  14 steps instead of 21: same bits, no faster)
#pragma unroll
    for (int j = 0; j < RPT; j++) {
        if (valid && rm.ok[j]) {
#pragma unroll
            for (int c = 0; c < 14; c++) P.G[(size_t)k * GSET + c * S + rm.row[j]] = getp(aQ[j], c);   // :372,:379
        }
        if (rm.ok[j]) st14_v4(myP + rm.row[j] * PR + cD, aQ[j]);
    }
#pragma unroll
    for (int j = 0; j < RPT7; j++) {
        if (inner && rm7.ok[j]) {
#pragma unroll
            for (int c = 0; c < 7; c++) P.G[(size_t)(k - 1) * GSET + SS + c * C + rm7.row[j]] = aR[j][c];   // :375
        }
        if (rm7.ok[j]) {
            T2* rr = reinterpret_cast<T2*>(myP + (7 + rm7.row[j]) * PR + cL);
            rr[0] = f2(aR[j][0], aR[j][1]); rr[1] = f2(aR[j][2], aR[j][3]);
            rr[2] = f2(aR[j][4], aR[j][5]); rr[3] = f2(aR[j][6], T(0));
        }
    }
    __syncwarp();
    T ti[RPT];                                                     // Qt_k^-1 q_k  (:411-415, :260)
    {
        T2 x[7];
        ld14_v4(x, vq);
#pragma unroll
        for (int j = 0; j < RPT; j++) ti[j] = hsum(dot14_acc(aQ[j], x, splat(T(0))));
    }

    cluster_sync_all<CL>();
    if (CL > 1) {
        if (rank > 0) pull_pd<CF>(smem, 0, KPC, rank - 1);
        __syncthreads();
    }

    // ---- phase 1b: phi, theta, gamma, theta^-1 ----------------------------------------------
    T2 th[RPT][7];
    {
        T2 phi[RPT][7];
#pragma unroll
        for (int j = 0; j < RPT; j++)
#pragma unroll
            for (int q = 0; q < 7; q++) { phi[j][q] = splat(T(0)); th[j][q] = aQ[j][q]; }   // + Qt_k^-1 (:466)
        const T* Qm = PW + (sl - 1) * PK + cD;                     // Qt_{k-1}^-1 rows
#pragma unroll
        for (int m = 0; m < 14; m++) {                                 // phi = A Qt_{k-1}^-1   (:398)
            T2 b[7];
            ld14_v4(b, Qm + m * PR);
            T a[RPT];
#pragma unroll
            for (int j = 0; j < RPT; j++) a[j] = myT[m * S + rm.row[j]];
            axpy_rows<RPT>(phi, a, b);
        }
        T br[RPT][7];
#pragma unroll
        for (int j = 0; j < RPT; j++)
#pragma unroll
            for (int c = 0; c < 7; c++) br[j][c] = T(0);
#pragma unroll
        for (int m = 0; m < 7; m++) {                                  // BR = B Rt_{k-1}^-1    (:406)
            const T2* rr = reinterpret_cast<const T2*>(myP + (7 + m) * PR + cL);
            const T2 u0 = rr[0], u1 = rr[1], u2 = rr[2], u3 = rr[3];
            const T rmv[7] = {u0.x, u0.y, u1.x, u1.y, u2.x, u2.y, u3.x};
#pragma unroll
            for (int j = 0; j < RPT; j++) {
                const T b = myP[m * PR + cL + rm.row[j]];
#pragma unroll
                for (int c = 0; c < 7; c++) br[j][c] = fm(b, rmv[c], br[j][c]);
            }
        }
#pragma unroll
        for (int m = 0; m < 14; m++) {                                 // + phi A^T             (:447)
            T2 b[7];
            ld14_v2(b, myT + m * S);                                   // column m of A
            T a[RPT];
#pragma unroll
            for (int j = 0; j < RPT; j++) a[j] = getp(phi[j], m);
            axpy_rows<RPT>(th, a, b);
        }
#pragma unroll
        for (int m = 0; m < 7; m++) {                                  // + BR B^T              (:473,:485)
            T2 b[7];
            ld14_v2(b, myP + m * PR + cL);                             // column m of B
            T a[RPT];
#pragma unroll
            for (int j = 0; j < RPT; j++) a[j] = br[j][m];
            axpy_rows<RPT>(th, a, b);
        }
        // gamma  (:416-443)
        {
            T2 x[7];
            ld14_v4(x, vqm);
            T rv[8];
            ld8(rv, vrm);
#pragma unroll
            for (int j = 0; j < RPT; j++) {
                const T s = hsum(dot14_acc(phi[j], x, splat(T(0))));
                T d = T(0);
#pragma unroll
                for (int c = 0; c < 7; c++) d = fm(br[j][c], rv[c], d);
                const T gm = (ti[j] - vc[rm.row[j]]) + (d + s);
                R.gamma[j] = inner ? -gm : -ti[j];                     // :531 / :274
            }
        }
#pragma unroll
        for (int j = 0; j < RPT; j++)
#pragma unroll
            for (int q = 0; q < 7; q++) R.Sd[j][q] = inner ? neg2(th[j][q]) : neg2(aQ[j][q]);   // :501 / :249
        // phi rows replace A in this knot's tile (only this knot's lanes read A): S[k][0] = -phi (:491)
        __syncwarp();
        if (k != 0) {
#pragma unroll
            for (int j = 0; j < RPT; j++) if (rm.ok[j]) st14_v2(myT + rm.row[j] * S, phi[j]);
        }
    }
    gj14<LPK, RPT>(th, t, piv);                                        // theta_k^-1            (:511-513)

    cluster_sync_all<CL>();                        // every knot (and every CTA's halo pull) is done with Qt^-1
    // Pd rows: -theta_k^-1 (:518-524), knot 0: -Qt_0 (:202-210)
#pragma unroll
    for (int j = 0; j < RPT; j++) {
        if (k == 0) ld14_v2(th[j], myT + rm.row[j] * S);
        if (rm.ok[j]) {
            T2 n[7];
#pragma unroll
            for (int q = 0; q < 7; q++) n[q] = neg2(th[j][q]);
            st14_v4(myP + rm.row[j] * PR + cD, n);
        }
    }
    cluster_sync_all<CL>();
    if (CL > 1) {
        if (rank > 0) pull_pd<CF>(smem, 0, KPC, rank - 1);
        if (rank + 1 < (unsigned)CL) pull_floats<CF>(TA + (KPC + 1) * TK, TA + 1 * TK, TK, rank + 1);
        __syncthreads();
    }

    // ---- phase 2: Pinv left block = theta_k^-1 phi_k theta_{k-1}^-1 (stored sign, :97-115) ----
    {
        T2 t1[RPT][7], pl[RPT][7];
#pragma unroll
        for (int j = 0; j < RPT; j++)
#pragma unroll
            for (int q = 0; q < 7; q++) { t1[j][q] = splat(T(0)); pl[j][q] = splat(T(0)); }
#pragma unroll
        for (int m = 0; m < 14; m++) {                                 // theta_k^-1 phi_k        (:100)
            T2 b[7];
            ld14_v2(b, myT + m * S);
            T a[RPT];
#pragma unroll
            for (int j = 0; j < RPT; j++) a[j] = getp(th[j], m);
            axpy_rows<RPT>(t1, a, b);
        }
        const T* Pm = PW + (sl - 1) * PK + cD;                     // -theta_{k-1}^-1 rows
#pragma unroll
        for (int m = 0; m < 14; m++) {                                 // ... theta_{k-1}^-1      (:102)
            T2 b[7];
            ld14_v4(b, Pm + m * PR);
            T a[RPT];
#pragma unroll
            for (int j = 0; j < RPT; j++) a[j] = -getp(t1[j], m);
            axpy_rows<RPT>(pl, a, b);
        }
        // own rows -> Pl columns; transposed -> Pr columns of the left neighbour (:118-136 by symmetry)
        T* left = PW + (sl - 1) * PK + cR;
        if (CL > 1 && g == 0 && rank > 0) left = map_rank<CL>(PW + KPC * PK + cR, rank - 1);
#pragma unroll
        for (int j = 0; j < RPT; j++) {
            if (!inner) {
#pragma unroll
                for (int q = 0; q < 7; q++) pl[j][q] = splat(T(0));
            }
            if (rm.ok[j]) {
                st14_v2(myP + rm.row[j] * PR + cL, pl[j]);
#pragma unroll
                for (int c = 0; c < 14; c++) left[c * PR + rm.row[j]] = getp(pl[j], c);
            }
        }
    }
    // last valid knot (and padding knots): no right neighbour writes their Pr columns
    if (!has_right) {
#pragma unroll
        for (int j = 0; j < RPT; j++)
            if (rm.ok[j]) {
#pragma unroll
                for (int c = 0; c < 14; c++) myP[c * PR + cR + rm.row[j]] = T(0);
            }
    }
    // ---- S left rows (own phi rows) and right rows (-phi_{k+1}^T, :537-557) ------------------
    {
        const T* phin = TA + (sl + 1) * TK;
#pragma unroll
        for (int j = 0; j < RPT; j++) {
            T2 a[7];
            ld14_v2(a, myT + rm.row[j] * S);
            const bool on = valid && rm.ok[j];
#pragma unroll
            for (int q = 0; q < 7; q++) R.Sl[j][q] = (inner && rm.ok[j]) ? neg2(a[q]) : splat(T(0));
#pragma unroll
            for (int c = 0; c < 14; c++) setp(R.Sr[j], c, (has_right && rm.ok[j]) ? -phin[c * S + rm.row[j]] : T(0));
            if (!on) {
#pragma unroll
                for (int q = 0; q < 7; q++) R.Sd[j][q] = splat(T(0));
                R.gamma[j] = T(0);
            }
        }
    }
    cluster_sync_all<CL>();                        // PROW complete (incl. rows scattered by neighbours)
    if (!valid) {                                  // padding knots: zero Pinv rows (their Pd holds 1/rho junk)
#pragma unroll
        for (int j = 0; j < RPT; j++)
            if (rm.ok[j]) {
#pragma unroll
                for (int q = 0; q < PR / 2; q++) reinterpret_cast<T2*>(myP + rm.row[j] * PR)[q] = splat(T(0));
            }
    }
}

// ---------------------------------------------------------------------------------------
// Cluster-wide sum with a bit-identical result in every thread of every CTA.
//  CL == 1: warp shuffle, one smem store per warp, __syncthreads, every thread adds the same
//           NW partials in the same order.
//  CL  > 1: every warp sends its partial to every CTA of the cluster with st.async (remote
//           store + complete_tx on the destination's mbarrier); each CTA waits on its OWN
//           mbarrier for NRED*4 (+ halo) bytes -- one-way DSMEM latency, no cluster barrier.
//           All lanes then combine the same NRED partials with the same butterfly.
// `mb` points at the two mbarriers (one per reduction kind), `par` at their phase parities.
// ---------------------------------------------------------------------------------------
// the NRED partials at `r`, added in four interleaved chains in a fixed order (broadcast loads)
template <int NRED, int NREDP>
__device__ __forceinline__ float sum_partials(const float* r) {
    const float4* r4 = reinterpret_cast<const float4*>(r);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int q = 0; q < NREDP / 4; q++) {
        const float4 v = r4[q];
        acc.x += v.x;
        if (4 * q + 1 < NRED) acc.y += v.y;
        if (4 * q + 2 < NRED) acc.z += v.z;
        if (4 * q + 3 < NRED) acc.w += v.w;
    }
    return (acc.x + acc.y) + (acc.z + acc.w);
}
__device__ __forceinline__ float sum4(const float* r) { const float4 v = *reinterpret_cast<const float4*>(r); return (v.x + v.y) + (v.z + v.w); }
__device__ __forceinline__ double sum4(const double* r) {
    const double2 a = reinterpret_cast<const double2*>(r)[0], b = reinterpret_cast<const double2*>(r)[1];
    return (a.x + a.y) + (b.x + b.y);
}
template <int NRED, int NREDP>
__device__ __forceinline__ double sum_partials(const double* r) {
    const double2* r2 = reinterpret_cast<const double2*>(r);
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
    for (int q = 0; q < NREDP / 2; q++) {
        const double2 v = r2[q];
        if (2 * q < NRED) acc[(2 * q) & 3] += v.x;
        if (2 * q + 1 < NRED) acc[(2 * q + 1) & 3] += v.y;
    }
    return (acc[0] + acc[1]) + (acc[2] + acc[3]);
}

// NRED (a, b) pairs at `r`, added in two interleaved chains per component in a fixed order (broadcast loads)
template <int NRED>
__device__ __forceinline__ float2 sum_pairs(const float2* r) {
    float2 a0 = make_float2(0.f, 0.f), a1 = make_float2(0.f, 0.f);
    if constexpr (NRED % 2 == 0 && NRED <= 64) {
        const float4* r4 = reinterpret_cast<const float4*>(r);
#pragma unroll
        for (int q = 0; q < NRED / 2; q++) {
            const float4 v = r4[q];
            a0.x += v.x; a0.y += v.y; a1.x += v.z; a1.y += v.w;
        }
    } else if constexpr (NRED <= 64) {
#pragma unroll
        for (int q = 0; q < NRED; q++) { const float2 v = r[q]; if (q & 1) { a1.x += v.x; a1.y += v.y; } else { a0.x += v.x; a0.y += v.y; } }
    } else {      // many partials: four pairs per lane, then a butterfly
        static_assert(NRED % 4 == 0 && NRED <= 128, "four pairs per lane");
        const int lane = threadIdx.x & 31;
        if (lane < NRED / 4) {
            const float4 u = reinterpret_cast<const float4*>(r)[2 * lane], v = reinterpret_cast<const float4*>(r)[2 * lane + 1];
            a0 = make_float2((u.x + u.z) + (v.x + v.z), (u.y + u.w) + (v.y + v.w));
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { a0.x += __shfl_xor_sync(0xffffffffu, a0.x, o); a0.y += __shfl_xor_sync(0xffffffffu, a0.y, o); }
    }
    return make_float2(a0.x + a1.x, a0.y + a1.y);
}
template <int NRED>
__device__ __forceinline__ double2 sum_pairs(const double2* r) {
    double2 a0 = make_double2(0.0, 0.0), a1 = make_double2(0.0, 0.0);
    if constexpr (NRED <= 64) {
#pragma unroll
        for (int q = 0; q < NRED; q++) { const double2 v = r[q]; if (q & 1) { a1.x += v.x; a1.y += v.y; } else { a0.x += v.x; a0.y += v.y; } }
    } else {
        static_assert(NRED % 4 == 0 && NRED <= 128, "four pairs per lane");
        const int lane = threadIdx.x & 31;
        if (lane < NRED / 4) {
            const double2 p0 = r[4 * lane], p1 = r[4 * lane + 1], p2 = r[4 * lane + 2], p3 = r[4 * lane + 3];
            a0 = make_double2((p0.x + p1.x) + (p2.x + p3.x), (p0.y + p1.y) + (p2.y + p3.y));
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { a0.x += __shfl_xor_sync(0xffffffffu, a0.x, o); a0.y += __shfl_xor_sync(0xffffffffu, a0.y, o); }
    }
    return make_double2(a0.x + a1.x, a0.y + a1.y);
}
// one (a, b) pair to a peer CTA's shared memory, completing its bytes on that CTA's mbarrier
__device__ __forceinline__ void st_async_pair(uint32_t a, float x, float y, uint32_t mb) {
    const unsigned long long v = (unsigned long long)__float_as_uint(x) | ((unsigned long long)__float_as_uint(y) << 32);
    asm volatile("st.async.shared::cluster.mbarrier::complete_tx::bytes.b64 [%0], %1, [%2];" ::"r"(a), "l"(v), "r"(mb) : "memory");
}
__device__ __forceinline__ void st_async_pair(uint32_t a, double x, double y, uint32_t mb) { st_async_f64(a, x, mb); st_async_f64(a + 8u, y, mb); }

// RADIX: warp sums as radix-8 + radix-4 exchanges (common.cuh).  Measured: faster in the single-GPU kernels (4096 x N=64
// fused -1.7 %, N=128 cluster PCG -2.2 %), SLOWER in the knot-split kernel (2 GPUs, N=512: 7.7 -> 8.6 us per iteration,
// profiles/r2o_knotsplit_ab.txt), which therefore keeps the five-level butterfly.
template <class CF, bool RADIX = true>
__device__ __forceinline__ typename CF::T cluster_sum(typename CF::T* red, uint64_t* mb, unsigned (&par)[2], int buf,
                                                      typename CF::T val, unsigned rank, unsigned halo_bytes) {
    using T = typename CF::T;
    constexpr int CL = CF::CL, NW = CF::NW, NRED = CF::NRED;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const T w = RADIX ? warp_sum(val) : warp_sum_butterfly(val);
    T* slot = red + buf * CF::NREDP + rank * NW + warp;
    if constexpr (CL == 1) {
        if (lane == 0) *slot = w;
        __syncthreads();
        return sum_partials<NRED, CF::NREDP>(red + buf * CF::NREDP);
    } else {
        if (threadIdx.x == 0) mbar_arrive_expect_tx(mb + buf, (unsigned)NRED * (unsigned)sizeof(T) + halo_bytes);
        if (lane < CL) st_async(mapa_u32(smem_u32(slot), (unsigned)lane), w, mapa_u32(smem_u32(mb + buf), (unsigned)lane));
        mbar_wait(mb + buf, par[buf]);
        par[buf] ^= 1u;
        if constexpr (NRED <= 64) {
            // every lane adds all NRED partials itself (broadcast LDS.128, four interleaved chains, fixed
            // order -> the same bits in every lane of every CTA): shorter than a second 5-level shuffle
            // tree (measured N=128, CL=8: 1.74 -> 1.35 us per iteration)
            return sum_partials<NRED, CF::NREDP>(red + buf * CF::NREDP);
        } else {
            // many partials (CL = 16 with 8 warps): four per lane, then the butterfly
            static_assert(NRED % 4 == 0 && NRED <= 128, "four partials per lane");
            T acc = T(0);
            if (lane < NRED / 4) {
                acc = sum4(red + buf * CF::NREDP + 4 * lane);
            }
            return RADIX ? warp_sum(acc) : warp_sum_butterfly(acc);
        }
    }
}

// ---------------------------------------------------------------------------------------
// Knot-range split across GPUs (BASELINE.json configs[3]): every rank (one GPU, one cluster)
// owns a contiguous knot range; the block-tridiagonal coupling needs, per PCG iteration, the
// boundary rows of S p / Pinv r from the two neighbour ranks and two scalar all-reduces.
// Both travel as {value, tag} 8-byte system-scope stores into a small MAILBOX in the
// destination GPU's memory (NVLink peer stores) and are consumed by polling the LOCAL mailbox:
// no host round trip, no NCCL call, no kernel boundary per iteration.  Tags are
// epoch + exchange counter, so a mailbox never needs clearing between solves.
// Mailbox layout (8-byte units): [2 solve parities] x ( red[2 bufs][16 ranks] | hal[5 phases][2 sides][16 rows] ).
// Consecutive solves use alternating halves: a rank that is already in solve j+1 (it only needs its two
// neighbours to get there) posts into the other half, so it can never overwrite a slot of solve j that a lagging
// warp of a third rank has not read yet; half j is reused by solve j+2, which no rank can reach before every rank
// has left solve j.  {value, tag} is ONE 64-bit store / load (single-copy atomic), never two words.
// ---------------------------------------------------------------------------------------
constexpr int SPLIT_MAXW = 16;
constexpr int SPLIT_HAL0 = 2 * SPLIT_MAXW;
constexpr int SPLIT_HALF_U2 = SPLIT_HAL0 + 5 * 2 * 16;      // one parity: 192 x 8 B
constexpr int SPLIT_MAIL_U2 = 2 * SPLIT_HALF_U2;            // 3 KB per rank
struct SplitCtx {
    unsigned grank, world;     // this cluster's rank among the knot-range owners
    unsigned epoch;            // tag base of this solve (identical on all ranks, above all earlier tags)
    uint2* const* mail;        // [world] mailbox of every rank (local or peer-mapped device pointers)
    unsigned parity;           // which half of every mailbox this solve uses
    unsigned long long timeout_ns;   // how long a poll may wait for a peer (%globaltimer)
    unsigned* diag;            // optional host-visible record of a timed-out poll: {1, rank, expected tag, seen tag}
};
__device__ __forceinline__ void sys_store_tagged(uint2* dst, float v, unsigned tag) {
    const unsigned long long w = (unsigned long long)__float_as_uint(v) | ((unsigned long long)tag << 32);
    asm volatile("st.relaxed.sys.global.b64 [%0], %1;" ::"l"(dst), "l"(w) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
// A dead peer must not hang the GPU: a poll that waits longer than X.timeout_ns (wall clock, configurable through
// MPCGPU_B200_KNOTSPLIT_TIMEOUT_MS, default 30 s -- a peer that is merely late, e.g. lazy module loading, is not
// dead) records what it was waiting for in host-visible memory and traps: a loud CUDA error on this rank.
static __device__ __noinline__ void sys_poll_timed_out(const SplitCtx& X, unsigned tag, unsigned seen) {
    if (X.diag) {
        X.diag[1] = X.grank; X.diag[2] = tag; X.diag[3] = seen;
        __threadfence_system();
        X.diag[0] = 1u;
        __threadfence_system();
    }
    __trap();
}
__device__ __forceinline__ float sys_poll_tagged(const SplitCtx& X, const uint2* src, unsigned tag) {
    unsigned long long w, t0 = 0;
    for (unsigned spin = 0;; spin++) {
        asm volatile("ld.relaxed.sys.global.b64 %0, [%1];" : "=l"(w) : "l"(src) : "memory");
        if ((unsigned)(w >> 32) == tag) return __uint_as_float((unsigned)w);
        if ((spin & 1023u) == 1023u) {                       // look at the clock every 1024 polls
            const unsigned long long now = globaltimer_ns();
            if (t0 == 0) t0 = now;
            else if (now - t0 > X.timeout_ns) sys_poll_timed_out(X, tag, (unsigned)(w >> 32));
        }
    }
}

// ---------------------------------------------------------------------------------------
// Sum of the NT per-thread partials a CTA has stored at `part` (after a CTA barrier), formed redundantly by every warp
// in a fixed order -- the same bits in every thread -- and cut into 7 steps so that a caller can spread it over the
// 7 column-pair steps of a matvec: the loads, the lane's own tree and the 5 butterfly levels then sit between
// independent FFMA2s instead of in front of a waiting CTA.
// ---------------------------------------------------------------------------------------
// Column pair w = 0..6 of a 14-vector slot in shared memory (16 B aligned), asked for in ascending order.  WIDE: float slots
// are fetched 16 B at a time (pairs w, w+1 together: 4 loads per block instead of 7).  That pays where the matvecs are
// issue-bound (two or more warps per scheduler: 4096 x N=64 fused 3.16 -> 3.10 ms) and costs where they are latency-bound
// (the 128-thread cluster tilings, one warp per scheduler: N=128 single PCG 89.4 -> 93.0 us), so the tiling chooses.
template <class T2, bool WIDE> struct XVec;
template <bool WIDE> struct XVec<float2, WIDE> {
    const float* v; float2 odd;
    __device__ __forceinline__ explicit XVec(const float* p) : v(p), odd(make_float2(0.f, 0.f)) {}
    __device__ __forceinline__ float2 get(int w) {
        if (!WIDE || w == 6) return reinterpret_cast<const float2*>(v)[w];
        if (w & 1) return odd;
        const float4 q = reinterpret_cast<const float4*>(v)[w >> 1];
        odd = make_float2(q.z, q.w);
        return make_float2(q.x, q.y);
    }
};
template <bool WIDE> struct XVec<double2, WIDE> {
    const double2* v;
    __device__ __forceinline__ explicit XVec(const double* p) : v(reinterpret_cast<const double2*>(p)) {}
    __device__ __forceinline__ double2 get(int w) { return v[w]; }
};

__device__ __forceinline__ unsigned bits_of(float x) { return __float_as_uint(x); }
__device__ __forceinline__ unsigned bits_of(double x) { return (unsigned)__double2hiint(x); }
__device__ __forceinline__ float or_bits(float x, unsigned m) { return __uint_as_float(__float_as_uint(x) | m); }
__device__ __forceinline__ double or_bits(double x, unsigned m) { return __hiloint2double(__double2hiint(x) | (int)m, __double2loint(x)); }
template <class T, int NT>
struct AllSum {
    static constexpr int V4 = 16 / sizeof(T);              // partials per 16 B load
    static constexpr int NQ = (NT / V4 + 31) / 32;         // 16 B loads per lane
    const T* part;
    T q[NQ][V4];
    T a, t, t2, t3;
    T den, quot;                                           // quot = sum / den, formed in the last step (the caller's next scalar)
    // The steps must STAY spread over the caller's instruction stream: left alone, ptxas pulls the whole chain to the front
    // of the matvec, where every level again waits out its shuffle with nothing to issue (measured: no gain).  Each step
    // therefore ORs (zero & dep) into one of its inputs, `dep` being bits of the caller's accumulator at that point and
    // `zero` a run-time zero the compiler cannot see through (blockIdx.z of a 1-D grid): a true data dependence on the
    // matvec's progress, one LOP3 per step, no change of value.
    unsigned zero;
    __device__ __forceinline__ AllSum(const T* p, T d) : part(p), a(T(0)), t(T(0)), t2(T(0)), t3(T(0)), den(d), quot(T(0)), zero(blockIdx.z) {}
    __device__ __forceinline__ void step(int w, unsigned dep) {
        const int lane = threadIdx.x & 31;
        const unsigned m = zero & dep;
        if (w == 0) {
#pragma unroll
            for (int i = 0; i < NQ; i++) {
                const int idx = lane + 32 * i;
                if (NT / V4 >= 32 * (i + 1) || idx < NT / V4) {
                    if constexpr (sizeof(T) == 4) {
                        const float4 v = reinterpret_cast<const float4*>(part)[idx];
                        q[i][0] = v.x; q[i][1] = v.y; q[i][2] = v.z; q[i][3] = v.w;
                    } else {
                        const double2 v = reinterpret_cast<const double2*>(part)[idx];
                        q[i][0] = v.x; q[i][1] = v.y;
                    }
                } else {
#pragma unroll
                    for (int e = 0; e < V4; e++) q[i][e] = T(0);
                }
            }
        } else if (w == 2) {          // (step 1 stays empty: 8 warps x 2 x 512 B of partials need ~64 cycles of the load pipe;
                                      //  loading them a step later and doubling up the exchanges instead was slower)
            T h[NQ];
#pragma unroll
            for (int i = 0; i < NQ; i++) {
                if constexpr (sizeof(T) == 4) h[i] = (or_bits(q[i][0], m) + q[i][1]) + (q[i][2] + q[i][3]);
                else h[i] = or_bits(q[i][0], m) + q[i][1];
            }
#pragma unroll
            for (int st = 1; st < NQ; st *= 2)
#pragma unroll
                for (int i = 0; i + st < NQ; i += 2 * st) h[i] += h[i + st];
            a = h[0];
            t = __shfl_xor_sync(0xffffffffu, a, 16);           // lane bits 4,3 in one radix-4 exchange
            t2 = __shfl_xor_sync(0xffffffffu, a, 8);
            t3 = __shfl_xor_sync(0xffffffffu, a, 24);
        } else if (w == 3) {
            a = (a + t2) + (or_bits(t, m) + t3);
            t = __shfl_xor_sync(0xffffffffu, a, 4);
        } else if (w <= 5) {
            a += or_bits(t, m);
            t = __shfl_xor_sync(0xffffffffu, a, w == 4 ? 2 : 1);
        } else if (w == 6) {
            a += or_bits(t, m);
            quot = divT(a, den);
        }
    }
};

// ---------------------------------------------------------------------------------------
// pcg_rows: preconditioned CG on the stored system S lambda = gamma.  Semantics (warm start,
// |r.Pinv r| < tol absolute exit, iteration count, max-iter flag) follow SURVEY 0.2 /
// oracle_pcg.  Per iteration: 2 reductions (__syncthreads for CL == 1; st.async + mbarrier
// across a cluster, no cluster-wide barrier) + 2 CTA barriers; the boundary rows of S p and Pinv r ride
// along with the partial sums, so neighbouring CTAs update their halo copies of r and p
// locally and need no further exchange.
// ---------------------------------------------------------------------------------------
template <class CF, bool SPLIT = false, int VAR = MPCB200_PCG_VARIANT>
__device__ __forceinline__ void pcg_rows(typename CF::T* smem, const RowsOf<CF>& R, typename CF::T* lam_g, int N,
                                         unsigned max_iter, typename CF::T exit_tol, unsigned* d_iters,
                                         uint8_t* d_exit, const SplitCtx* X = nullptr) {
    using T = typename CF::T;
    using T2 = typename CF::T2;
    static_assert(!SPLIT || CF::F32, "the knot-range split mailboxes carry fp32 values");
    constexpr int S = 14, KPC = CF::KPC, CL = CF::CL, PR = CF::PR, PK = CF::PK, VK = CF::VK;
    constexpr int LPK = CF::LPK, RPT = CF::RPT;
    const int tid = threadIdx.x;
    const int g = tid / LPK, t = tid % LPK;
    const unsigned rank = cluster_rank<CL>();
    const int k0 = rank * KPC, k = k0 + g;
    const bool valid = k < N;
    const RowMap<LPK, RPT, 14> rm(t);
    const bool first_g = (g == 0), last_g = (g == KPC - 1);
    const bool has_l = rank > 0, has_r = rank + 1 < (unsigned)CL;
    // knot-range split: neighbours of the cluster's outer CTAs live on other GPUs
    const bool gl = SPLIT && rank == 0 && X->grank > 0;
    const bool gr = SPLIT && rank + 1 == (unsigned)CL && X->grank + 1 < X->world;
    unsigned gseq[2] = {SPLIT ? X->epoch : 0u, SPLIT ? X->epoch : 0u};
    const int sl = g + 1;

    T* pv = smem + CF::oPv;
    T* rv = smem + CF::oRv;
    T* hal = smem + CF::oHal;   // [phase][side][16]; side 0: pushed by my LEFT neighbour CTA
    T* red = smem + CF::oRed;
    uint64_t* mb = reinterpret_cast<uint64_t*>(smem + CF::oMb);
    unsigned par[2] = {0u, 0u};
    const unsigned hb = (CL > 1) ? ((has_l ? 14u * (unsigned)sizeof(T) : 0u) + (has_r ? 14u * (unsigned)sizeof(T) : 0u)) : 0u;   // halo bytes per phase
    if (CL > 1 && tid == 0) { mbar_init(mb, 1); mbar_init(mb + 1, 1); fence_mbar_init_cluster(); }
    const T* myP = smem + CF::oP + sl * PK;
    constexpr bool TM = CF::TM;
    uint32_t* tm_word = reinterpret_cast<uint32_t*>(smem + CF::oTm);
    uint32_t tcol = 0;      // TMEM address of this lane's first column
    if (TM) {               // published by the barriers below, long before the first use
        if (tid < 32) tmem_alloc(tm_word, CF::TCOLS);
        tmem_fence_before_sync();
    }

    // The two block-row matvecs of an iteration.  ONE accumulation order everywhere (registers, shared memory, tensor
    // memory, float or double), so every storage variant of a tiling gives the same bits: per row a single pair-wise
    // accumulator runs over the column pairs w = 0..6 and, for each w, over the blocks left -> diagonal -> right
    // (the oracle's bdmv is the same sum row by row); the two halves of the pair are added at the end.  One chain per
    // row instead of one per block saves two of the three horizontal sums (round 1: 5 FADD per row and matvec; the
    // FP32 pipe, not the issue slots, bounds this loop -- see DESIGN.md): 4 rows per lane and 2 warps per scheduler
    // still cover the FFMA2 latency.
    //
    // Tensor-memory tilings (TM): registers hold Sl, Sd (R) and Pd (below); tensor memory holds Sr | Pl | Pr as 7 chunks
    // of 2*RPT columns (column pair w of my RPT rows of Sr) followed by 7 chunks of 4*RPT (Pl pair w | Pr pair w),
    // streamed one chunk ahead of the FFMA2s (tcgen05.wait::ld covers every load in flight: wait, THEN start the next).
    // The tensor-memory read path delivers ~1/4 of what the FFMA2 pipe can consume: with one block of S and two of Pinv
    // there (round 1: all of Pinv, none of S) neither matvec is bound by it.
    T2 Pd[TM ? RPT : 1][7];
    auto s_mv_h = [&](const T* vec, T (&y)[RPT], AllSum<T, CF::NT>* hook) {   // hook->step(w): independent work spread over the 7 steps
        XVec<T2, (CF::NT > 128)> vl(vec + (sl - 1) * VK), vd(vec + sl * VK), vr(vec + (sl + 1) * VK);
        T2 acc[RPT];
#pragma unroll
        for (int j = 0; j < RPT; j++) acc[j] = splat(T(0));
        if constexpr (TM) {
            TmemChunk<2 * RPT> buf[2];
            buf[0].ld(tcol);
#pragma unroll
            for (int w = 0; w < 7; w++) {
                if (hook) hook->step(w, bits_of(acc[0].x));      // acc[0] as step w - 1 left it: long since complete
                const T2 xl = vl.get(w), xd = vd.get(w), xr = vr.get(w);
#pragma unroll
                for (int j = 0; j < RPT; j++) { acc[j] = fma2(R.Sl[j][w], xl, acc[j]); acc[j] = fma2(R.Sd[j][w], xd, acc[j]); }
                buf[w & 1].wait();
                if (w + 1 < 7) buf[(w + 1) & 1].ld(tcol + (w + 1) * 2 * RPT);
#pragma unroll
                for (int j = 0; j < RPT; j++)
                    acc[j] = fma2(f2(__uint_as_float(buf[w & 1].q[2 * j]), __uint_as_float(buf[w & 1].q[2 * j + 1])), xr, acc[j]);
            }
        } else {
#pragma unroll
            for (int w = 0; w < 7; w++) {
                if (hook) hook->step(w, bits_of(acc[0].x));
                const T2 xl = vl.get(w), xd = vd.get(w), xr = vr.get(w);
#pragma unroll
                for (int j = 0; j < RPT; j++) {
                    acc[j] = fma2(R.Sl[j][w], xl, acc[j]); acc[j] = fma2(R.Sd[j][w], xd, acc[j]); acc[j] = fma2(R.Sr[j][w], xr, acc[j]);
                }
            }
        }
#pragma unroll
        for (int j = 0; j < RPT; j++) y[j] = hsum(acc[j]);
    };
    auto s_mv = [&](const T* vec, T (&y)[RPT]) { s_mv_h(vec, y, nullptr); };
    auto pinv_mv = [&](const T* vec, T (&y)[RPT]) {
        XVec<T2, (CF::NT > 128)> vl(vec + (sl - 1) * VK), vd(vec + sl * VK), vr(vec + (sl + 1) * VK);
        T2 acc[RPT];
#pragma unroll
        for (int j = 0; j < RPT; j++) acc[j] = splat(T(0));
        if constexpr (TM) {
            TmemChunk<4 * RPT> buf[2];
            const uint32_t base = tcol + 7 * 2 * RPT;
            buf[0].ld(base);
#pragma unroll
            for (int w = 0; w < 7; w++) {
                const T2 xl = vl.get(w), xd = vd.get(w), xr = vr.get(w);
                buf[w & 1].wait();
                if (w + 1 < 7) buf[(w + 1) & 1].ld(base + (w + 1) * 4 * RPT);
#pragma unroll
                for (int j = 0; j < RPT; j++) {
                    acc[j] = fma2(f2(__uint_as_float(buf[w & 1].q[2 * j]), __uint_as_float(buf[w & 1].q[2 * j + 1])), xl, acc[j]);
                    acc[j] = fma2(Pd[j][w], xd, acc[j]);
                    acc[j] = fma2(f2(__uint_as_float(buf[w & 1].q[2 * RPT + 2 * j]), __uint_as_float(buf[w & 1].q[2 * RPT + 2 * j + 1])), xr, acc[j]);
                }
            }
        } else {      // the Pinv rows from shared memory (row = [Pd | Pl | Pr | pad]), same order
#pragma unroll
            for (int w = 0; w < 7; w++) {
                const T2 xl = vl.get(w), xd = vd.get(w), xr = vr.get(w);
#pragma unroll
                for (int j = 0; j < RPT; j++) {
                    const T* row = myP + rm.row[j] * PR;
                    acc[j] = fma2(*reinterpret_cast<const T2*>(row + CF::cL + 2 * w), xl, acc[j]);
                    acc[j] = fma2(*reinterpret_cast<const T2*>(row + CF::cD + 2 * w), xd, acc[j]);
                    acc[j] = fma2(*reinterpret_cast<const T2*>(row + CF::cR + 2 * w), xr, acc[j]);
                }
            }
        }
#pragma unroll
        for (int j = 0; j < RPT; j++) y[j] = rm.ok[j] ? hsum(acc[j]) : T(0);
    };
    auto put = [&](T* dst, const T (&v)[RPT]) {        // my rows of a 16-T vector slot
#pragma unroll
        for (int j = 0; j < RPT; j++) if (rm.ok[j]) dst[rm.row[j]] = v[j];
    };
    // my rows -> a 16-T slot of a neighbour CTA, completing bytes on ITS mbarrier `buf`
    auto push = [&](T* local_dst, unsigned dst_rank, int buf, const T (&v)[RPT]) {
        const uint32_t ra = mapa_u32(smem_u32(local_dst), dst_rank), rb = mapa_u32(smem_u32(mb + buf), dst_rank);
#pragma unroll
        for (int j = 0; j < RPT; j++) if (rm.ok[j]) st_async(ra + (unsigned)sizeof(T) * rm.row[j], v[j], rb);
    };
    // my rows -> mailbox [phase][side] of rank `dst` (another GPU), tagged
    auto gpush = [&](int phase, int side, unsigned dst, unsigned tag, const T (&v)[RPT]) {
        uint2* m = X->mail[dst] + X->parity * SPLIT_HALF_U2 + SPLIT_HAL0 + (phase * 2 + side) * 16;
#pragma unroll
        for (int j = 0; j < RPT; j++) if (rm.ok[j]) sys_store_tagged(m + rm.row[j], v[j], tag);
    };
    auto gpull = [&](int phase, int side, int r, unsigned tag) {
        return sys_poll_tagged(*X, X->mail[X->grank] + X->parity * SPLIT_HALF_U2 + SPLIT_HAL0 + (phase * 2 + side) * 16 + r, tag);
    };
    // boundary rows of the neighbour RANKS -> my halo slots (pre-loop exchanges, phases 2..4)
    auto gpull_rows = [&](int phase, unsigned tag, T* left_slot, T* right_slot) {
#pragma unroll
        for (int j = 0; j < RPT; j++) {
            if (!rm.ok[j]) continue;
            const int r = rm.row[j];
            if (first_g && gl) left_slot[r] = gpull(phase, 0, r, tag);
            if (last_g && gr)  right_slot[r] = gpull(phase, 1, r, tag);
        }
    };
    // all-reduce of the per-rank totals (w is already identical in every thread of this cluster)
#ifndef MPCB200_SPLIT_POLL1
#define MPCB200_SPLIT_POLL1 1     // 1: ONE warp per CTA polls the mailbox and hands the sum to the others through shared memory
#endif
    auto gsum = [&](int buf, unsigned tag, T w) {
        if (rank == 0 && tid < (int)X->world) sys_store_tagged(X->mail[tid] + X->parity * SPLIT_HALF_U2 + buf * SPLIT_MAXW + X->grank, w, tag);
        const int lane = tid & 31;
        if constexpr (MPCB200_SPLIT_POLL1) {
            T* bc = smem + CF::oTm + 1 + buf;          // two spare words beside the tensor-memory base address
            if (tid < 32) {
                T v = T(0);
                if (lane < (int)X->world) v = sys_poll_tagged(*X, X->mail[X->grank] + X->parity * SPLIT_HALF_U2 + buf * SPLIT_MAXW + lane, tag);
                __syncwarp();
                v = warp_sum_butterfly(v);
                if (lane == 0) *bc = v;
            }
            __syncthreads();
            return *bc;
        } else {
            T v = T(0);
            if (lane < (int)X->world) v = sys_poll_tagged(*X, X->mail[X->grank] + X->parity * SPLIT_HALF_U2 + buf * SPLIT_MAXW + lane, tag);
            __syncwarp();
            return warp_sum_butterfly(v);
        }
    };
    auto dotp = [&](const T (&a)[RPT], const T (&b)[RPT]) {
        T d = T(0);
#pragma unroll
        for (int j = 0; j < RPT; j++) d = fm(a[j], b[j], d);   // invalid slots hold zeros
        return d;
    };

    T* hl_p = pv;                        // left halo slot
    T* hr_p = pv + (KPC + 1) * VK;       // right halo slot
    T* hl_r = rv;
    T* hr_r = rv + (KPC + 1) * VK;
    const T zero[RPT] = {};

    // ---- r = gamma - S lambda0 --------------------------------------------------------------
    T lam[RPT], rr[RPT], pp[RPT], rt[RPT], y[RPT];
#pragma unroll
    for (int j = 0; j < RPT; j++) lam[j] = (valid && rm.ok[j]) ? lam_g[(size_t)k * S + rm.row[j]] : T(0);
    put(pv + sl * VK, lam);
    if (first_g) {
        const bool ok = (k0 >= 1 && k0 - 1 < N);
#pragma unroll
        for (int j = 0; j < RPT; j++) if (rm.ok[j]) hl_p[rm.row[j]] = ok ? lam_g[(size_t)(k0 - 1) * S + rm.row[j]] : T(0);
    }
    if (last_g) {
        const bool ok = (k0 + KPC < N);
#pragma unroll
        for (int j = 0; j < RPT; j++) if (rm.ok[j]) hr_p[rm.row[j]] = ok ? lam_g[(size_t)(k0 + KPC) * S + rm.row[j]] : T(0);
    }
    const unsigned tg0 = SPLIT ? X->epoch + 1u : 0u;     // tag of the three pre-loop exchanges
    if (SPLIT) {                                         // warm-start rows of the neighbour ranks
        if (first_g && gl) gpush(2, 1, X->grank - 1, tg0, lam);
        if (last_g && gr)  gpush(2, 0, X->grank + 1, tg0, lam);
        gpull_rows(2, tg0, hl_p, hr_p);
    }
    __syncthreads();
    if constexpr (TM) {   // my PROW rows are final (schur_rows / load_rows ended on a barrier): Sr, Pl, Pr -> tensor memory, Pd -> registers
        tmem_fence_after_sync();
        const int warp = tid >> 5;
        tcol = *tm_word + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)((warp >> 2) * CF::TPT);
#pragma unroll
        for (int w = 0; w < 7; w++) {
            TmemChunk<2 * RPT> cs;
            TmemChunk<4 * RPT> cp;
#pragma unroll
            for (int j = 0; j < RPT; j++) {
                const T2 pl = *reinterpret_cast<const T2*>(myP + rm.row[j] * PR + CF::cL + 2 * w);
                const T2 pr = *reinterpret_cast<const T2*>(myP + rm.row[j] * PR + CF::cR + 2 * w);
                Pd[j][w] = *reinterpret_cast<const T2*>(myP + rm.row[j] * PR + CF::cD + 2 * w);
                cs.q[2 * j] = __float_as_uint(R.Sr[j][w].x); cs.q[2 * j + 1] = __float_as_uint(R.Sr[j][w].y);
                cp.q[2 * j] = __float_as_uint(pl.x); cp.q[2 * j + 1] = __float_as_uint(pl.y);
                cp.q[2 * RPT + 2 * j] = __float_as_uint(pr.x); cp.q[2 * RPT + 2 * j + 1] = __float_as_uint(pr.y);
            }
            cs.st(tcol + w * 2 * RPT);
            cp.st(tcol + 7 * 2 * RPT + w * 4 * RPT);
        }
        tmem_wait_st();
    }
    s_mv(pv, y);
#pragma unroll
    for (int j = 0; j < RPT; j++) rr[j] = rm.ok[j] ? R.gamma[j] - y[j] : T(0);
    put(rv + sl * VK, rr);
    if (first_g) { if (has_l) put(map_rank<CL>(hr_r, rank - 1), rr); else if (gl) gpush(3, 1, X->grank - 1, tg0, rr); else put(hl_r, zero); }
    if (last_g)  { if (has_r) put(map_rank<CL>(hl_r, rank + 1), rr); else if (gr) gpush(3, 0, X->grank + 1, tg0, rr); else put(hr_r, zero); }
    if (SPLIT) gpull_rows(3, tg0, hl_r, hr_r);
    cluster_sync_all<CL>();

    unsigned iter = 0;
    bool max_iter_exit = true;
    constexpr int V = (VAR == 2 && (CL > 1 || SPLIT)) ? 0 : VAR;    // variant 2 exists for the single-CTA tilings
    if constexpr (V == 2) {
    // ---- S p by recurrence: q = S p is never formed from p after the first iteration.  With rt = Pinv r published for
    // the matvec S rt anyway, every thread leaves its partial of r.rt next to it, ONE barrier covers both, and the
    // sum of the partials (AllSum) runs inside the matvec; then p <- rt + beta p, q <- S rt + beta q.  alpha, beta and the
    // exit test are those of the loop below (two inner products per iteration); one of the two reductions and one
    // barrier leave the critical path.  Checked against oracle_pcg_sprec; the parity criteria are asserted against the
    // reference's iteration (oracle_pcg) like every other variant.
    T* part = smem + CF::oT;                 // [NT] partials of the hidden reduction (the Schur scratch tiles are dead by now)
    T qq[RPT];
    pinv_mv(rv, rt);
#pragma unroll
    for (int j = 0; j < RPT; j++) pp[j] = rt[j];
    put(pv + sl * VK, pp);
    if (first_g) put(hl_p, zero);
    if (last_g) put(hr_p, zero);
    T eta = cluster_sum<CF>(red, mb, par, 1, dotp(rr, rt), rank, hb);      // its barrier publishes pv
    s_mv(pv, qq);
    for (; iter < max_iter; iter++) {
        const T v = cluster_sum<CF>(red, mb, par, 0, dotp(pp, qq), rank, hb);
        if (v == T(0)) { max_iter_exit = false; break; }
        const T alpha = divT(eta, v);
#pragma unroll
        for (int j = 0; j < RPT; j++) { lam[j] = fm(alpha, pp[j], lam[j]); rr[j] = fm(-alpha, qq[j], rr[j]); }
        put(rv + sl * VK, rr);
        __syncthreads();
        pinv_mv(rv, rt);
        put(pv + sl * VK, rt);
        part[tid] = dotp(rr, rt);
        __syncthreads();
#ifdef MPCB200_V2_EXPOSED      // A/B builds: the same recurrence with the reduction NOT overlapped
        const T eta_new = cluster_sum<CF>(red, mb, par, 1, part[tid], rank, hb);
        s_mv(pv, y);
        if (absT(eta_new) < exit_tol) { iter++; max_iter_exit = false; break; }
        const T beta = divT(eta_new, eta);
#else
        AllSum<T, CF::NT> sum(part, eta);
        s_mv_h(pv, y, &sum);
        const T eta_new = sum.a;
        if (absT(eta_new) < exit_tol) { iter++; max_iter_exit = false; break; }
        const T beta = sum.quot;
#endif
        eta = eta_new;
#pragma unroll
        for (int j = 0; j < RPT; j++) { pp[j] = fm(beta, pp[j], rt[j]); qq[j] = fm(beta, qq[j], y[j]); }
    }
    } else if constexpr (V == 1) {
    // ---- single-reduction PCG (Chronopoulos & Gear): u = Pinv r, w = S u, ONE reduction of the pair
    // (r.u, u.w) per iteration; p.Sp = u.w - beta^2 (p.Sp)_old; q = S p by recurrence.  Same Krylov
    // iterates as the loop below in exact arithmetic, same exit test / iteration count / flag
    // (checked against oracle_pcg_cg1).  Per iteration: 3 CTA barriers, one shuffle tree.
    static_assert(!SPLIT, "the knot-range split runs the two-reduction iteration");
    T* hw = hal;               // [side][VK] boundary rows of S u pushed by the neighbour CTAs (side 0: my LEFT one)
    T* hq = hal + 2 * VK;      // [side][VK] q rows of the neighbours' boundary knots, by the same recurrence
    T2* red2 = reinterpret_cast<T2*>(red);
    if (CL > 1) { if (first_g) put(hq, zero); if (last_g) put(hq + VK, zero); }
    T qq[RPT];
#pragma unroll
    for (int j = 0; j < RPT; j++) { pp[j] = T(0); qq[j] = T(0); }
    // rt = Pinv r -> published in pv (own rows + the neighbour CTAs' halo slots) ; y = S rt
    auto mv_pair = [&]() {
        pinv_mv(rv, rt);
        put(pv + sl * VK, rt);
        if constexpr (CL > 1) {
            if (tid == 0) mbar_arrive_expect_tx(mb + 0, hb);
            if (first_g && has_l) push(hr_p, rank - 1, 0, rt);
            if (last_g && has_r)  push(hl_p, rank + 1, 0, rt);
            __syncthreads();
            mbar_wait(mb + 0, par[0]);
            par[0] ^= 1u;
        } else {
            __syncthreads();
        }
        s_mv(pv, y);
    };
    // cluster-wide sums of (a, b), bit-identical in every thread; the boundary rows of y travel with them
    auto sum2 = [&](T a, T b) -> T2 {
        const int lane = tid & 31, warp = tid >> 5;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { a += __shfl_xor_sync(0xffffffffu, a, o); b += __shfl_xor_sync(0xffffffffu, b, o); }
        T2* slot = red2 + rank * CF::NW + warp;
        if constexpr (CL == 1) {
            if (lane == 0) *slot = f2(a, b);
            __syncthreads();
        } else {
            if (tid == 0) mbar_arrive_expect_tx(mb + 1, (unsigned)CF::NRED * (unsigned)sizeof(T2) + hb);
            if (lane < CL) st_async_pair(mapa_u32(smem_u32(slot), (unsigned)lane), a, b, mapa_u32(smem_u32(mb + 1), (unsigned)lane));
            if (first_g && has_l) push(hw + 1 * VK, rank - 1, 1, y);
            if (last_g && has_r)  push(hw + 0 * VK, rank + 1, 1, y);
            mbar_wait(mb + 1, par[1]);
            par[1] ^= 1u;
        }
        return sum_pairs<CF::NRED>(red2);
    };
    if (first_g && !has_l) put(hl_p, zero);
    if (last_g && !has_r)  put(hr_p, zero);
    mv_pair();
    T2 gd = sum2(dotp(rr, rt), dotp(rt, y));
    T gam_old = T(1), c_old = T(0);
    for (; iter < max_iter; iter++) {
        const T gam = gd.x;
        const T v = fm(-(gam * gam), c_old, gd.y);          // p.Sp
        if (v == T(0)) { max_iter_exit = false; break; }
        const T alpha = gam / v;
        const T beta = iter ? gam / gam_old : T(0);
#pragma unroll
        for (int j = 0; j < RPT; j++) {
            pp[j] = fm(beta, pp[j], rt[j]);
            qq[j] = fm(beta, qq[j], y[j]);
            lam[j] = fm(alpha, pp[j], lam[j]);
            rr[j] = fm(-alpha, qq[j], rr[j]);
        }
        put(rv + sl * VK, rr);
        if (CL > 1) {       // the neighbours' boundary knots: the same recurrences on the pushed rows of S u
#pragma unroll
            for (int j = 0; j < RPT; j++) {
                if (!rm.ok[j]) continue;
                const int r = rm.row[j];
                if (first_g && has_l) { const T q = fm(beta, hq[r], hw[r]); hq[r] = q; hl_r[r] = fm(-alpha, q, hl_r[r]); }
                if (last_g && has_r)  { const T q = fm(beta, hq[VK + r], hw[VK + r]); hq[VK + r] = q; hr_r[r] = fm(-alpha, q, hr_r[r]); }
            }
        }
        c_old = (v / gam) / gam;     // off the critical path: needed after the next reduction
        gam_old = gam;
        __syncthreads();
        mv_pair();
        gd = sum2(dotp(rr, rt), dotp(rt, y));
        if (absT(gd.x) < exit_tol) { iter++; max_iter_exit = false; break; }
    }
    } else {
    // ---- p = rt = Pinv r ; eta = r.rt ----------------------------------------------------------
    pinv_mv(rv, rt);
#pragma unroll
    for (int j = 0; j < RPT; j++) pp[j] = rt[j];
    put(pv + sl * VK, pp);                   // pv was last read before the barrier above
    if (first_g) { if (has_l) push(hr_p, rank - 1, 1, pp); else if (gl) gpush(4, 1, X->grank - 1, tg0, pp); else put(hl_p, zero); }
    if (last_g)  { if (has_r) push(hl_p, rank + 1, 1, pp); else if (gr) gpush(4, 0, X->grank + 1, tg0, pp); else put(hr_p, zero); }
    if (SPLIT) gpull_rows(4, tg0, hl_p, hr_p);
    T eta = cluster_sum<CF, !SPLIT>(red, mb, par, 1, dotp(rr, rt), rank, hb);
    if (SPLIT) eta = gsum(1, ++gseq[1], eta);
    // the p rows stored above are read by OTHER warps in the first S p: for CL > 1 the reduction ends on an
    // mbarrier wait, which orders the st.async bytes only, so the ordinary stores need a CTA barrier of their own
    // (CL == 1: cluster_sum already contains one); once per solve
    if (CL > 1 || SPLIT) __syncthreads();

    for (; iter < max_iter; iter++) {
        // ---- ups = S p ; v = p.ups -----------------------------------------------------------
        s_mv(pv, y);
        if (CL > 1) {
            if (first_g && has_l) push(hal + (0 * 2 + 1) * VK, rank - 1, 0, y);
            if (last_g && has_r)  push(hal + (0 * 2 + 0) * VK, rank + 1, 0, y);
        }
        if (SPLIT) {
            if (first_g && gl) gpush(0, 1, X->grank - 1, gseq[0] + 1, y);
            if (last_g && gr)  gpush(0, 0, X->grank + 1, gseq[0] + 1, y);
        }
        T v = cluster_sum<CF, !SPLIT>(red, mb, par, 0, dotp(pp, y), rank, hb);
        if (SPLIT) v = gsum(0, ++gseq[0], v);
        if (v == T(0)) { max_iter_exit = false; break; }
        const T alpha = divT(eta, v);
#pragma unroll
        for (int j = 0; j < RPT; j++) { lam[j] = fm(alpha, pp[j], lam[j]); rr[j] = fm(-alpha, y[j], rr[j]); }
        put(rv + sl * VK, rr);
        if (CL > 1) {
#pragma unroll
            for (int j = 0; j < RPT; j++) {
                if (!rm.ok[j]) continue;
                const int r = rm.row[j];
                if (first_g && has_l) hl_r[r] = fm(-alpha, hal[(0 * 2 + 0) * VK + r], hl_r[r]);
                if (last_g && has_r)  hr_r[r] = fm(-alpha, hal[(0 * 2 + 1) * VK + r], hr_r[r]);
            }
        }
        if (SPLIT) {
#pragma unroll
            for (int j = 0; j < RPT; j++) {
                if (!rm.ok[j]) continue;
                const int r = rm.row[j];
                if (first_g && gl) hl_r[r] = fm(-alpha, gpull(0, 0, r, gseq[0]), hl_r[r]);
                if (last_g && gr)  hr_r[r] = fm(-alpha, gpull(0, 1, r, gseq[0]), hr_r[r]);
            }
        }
        __syncthreads();
        // ---- rt = Pinv r ; eta_new = r.rt ----------------------------------------------------
        pinv_mv(rv, rt);
        if (CL > 1) {
            if (first_g && has_l) push(hal + (1 * 2 + 1) * VK, rank - 1, 1, rt);
            if (last_g && has_r)  push(hal + (1 * 2 + 0) * VK, rank + 1, 1, rt);
        }
        if (SPLIT) {
            if (first_g && gl) gpush(1, 1, X->grank - 1, gseq[1] + 1, rt);
            if (last_g && gr)  gpush(1, 0, X->grank + 1, gseq[1] + 1, rt);
        }
        T eta_new = cluster_sum<CF, !SPLIT>(red, mb, par, 1, dotp(rr, rt), rank, hb);
        if (SPLIT) eta_new = gsum(1, ++gseq[1], eta_new);
        if (absT(eta_new) < exit_tol) { iter++; max_iter_exit = false; break; }
        const T beta = divT(eta_new, eta);
        eta = eta_new;
#pragma unroll
        for (int j = 0; j < RPT; j++) pp[j] = fm(beta, pp[j], rt[j]);
        put(pv + sl * VK, pp);
        if (CL > 1) {
#pragma unroll
            for (int j = 0; j < RPT; j++) {
                if (!rm.ok[j]) continue;
                const int r = rm.row[j];
                if (first_g && has_l) hl_p[r] = fm(beta, hl_p[r], hal[(1 * 2 + 0) * VK + r]);
                if (last_g && has_r)  hr_p[r] = fm(beta, hr_p[r], hal[(1 * 2 + 1) * VK + r]);
            }
        }
        if (SPLIT) {
#pragma unroll
            for (int j = 0; j < RPT; j++) {
                if (!rm.ok[j]) continue;
                const int r = rm.row[j];
                if (first_g && gl) hl_p[r] = fm(beta, hl_p[r], gpull(1, 0, r, gseq[1]));
                if (last_g && gr)  hr_p[r] = fm(beta, hr_p[r], gpull(1, 1, r, gseq[1]));
            }
        }
        __syncthreads();
    }
    }   // VAR
#pragma unroll
    for (int j = 0; j < RPT; j++) if (valid && rm.ok[j]) lam_g[(size_t)k * S + rm.row[j]] = lam[j];
    if (rank == 0 && tid == 0) { *d_iters = iter; *d_exit = max_iter_exit ? 1 : 0; }
    // no CTA may exit while a peer can still address its shared memory
    if (CL > 1) cluster_sync_all<CL>();
    if (TM) {
        tmem_fence_before_sync();
        __syncthreads();
        if (tid < 32) tmem_dealloc(*tm_word, CF::TCOLS);
    }
}

}  // namespace mpcb200
