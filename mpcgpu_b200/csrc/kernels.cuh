// kernels.cuh -- __global__ entry points built from kkt_device.cuh.
//
// Grid: batch * CL CTAs, one thread-block cluster of CL CTAs per KKT instance, KPC knots per
// CTA, LPK lanes per knot.  Instance b's buffers start at base + b * <elements per instance>
// (the reference's single-instance buffers, concatenated).
#pragma once
#include "kkt_device.cuh"
#include "dz_device.cuh"

namespace mpcb200 {

template <class T>
struct BatchArgsT {
    int N;
    int batch;
    // per-instance element strides
    size_t sG, sC, sg, sc, sS, sv;
    T* G; const T* C; const T* g; const T* c;
    T rho;
    T* S; T* Pinv; T* gamma;   // two-stage outputs / inputs
    T* lambda;
    unsigned max_iter; T exit_tol;
    unsigned* iters; uint8_t* exits;
    T* dz = nullptr; size_t sdz = 0;   // optional compute_dz epilogue (fused) / output of dz_kernel
    const T* rho_dev = nullptr;        // optional: rho read from device memory (launches replayed from a CUDA graph)
    size_t rho_stride = 0;             // 0: one rho for the batch; 1: rho_dev[b] per instance (batched SQP iterations)
    int pf_dist = 0;                   // > 0: while instance b iterates, prefetch the inputs of instance b + pf_dist into L2
};
using BatchArgs = BatchArgsT<float>;

// PROW column offset of block column `blk` (0 left, 1 diag, 2 right) of the reference layout
template <class CF>
__device__ __forceinline__ int prow_col(int blk) { return blk == 1 ? CF::cD : (blk == 0 ? CF::cL : CF::cR); }

// store to the reference [N][3][S*S] column-major layout (store_block_bd semantics,
// include/pcg/linsys_setup_no_precond.cuh:29-38); corner blocks are written as zeros.
// S rows come from registers, Pinv rows from shared memory (coalesced).
template <class CF>
__device__ __forceinline__ void store_rows(const typename CF::T* smem, const RowsOf<CF>& R, typename CF::T* Sg,
                                           typename CF::T* Pg, typename CF::T* gam, int N) {
    using T = typename CF::T;
    constexpr int S = 14, SS = S * S;
    const int g = threadIdx.x / CF::LPK, t = threadIdx.x % CF::LPK;
    const int k0 = cluster_rank<CF::CL>() * CF::KPC;
    const int k = k0 + g;
    const RowMap<CF::LPK, CF::RPT, 14> rm(t);
    if (k < N) {
        T* s0 = Sg + (size_t)k * 3 * SS;
#pragma unroll
        for (int j = 0; j < CF::RPT; j++) {
            if (!rm.ok[j]) continue;
            const int r = rm.row[j];
#pragma unroll
            for (int c = 0; c < S; c++) {
                s0[0 * SS + c * S + r] = getp(R.Sl[j], c);
                s0[1 * SS + c * S + r] = getp(R.Sd[j], c);
                s0[2 * SS + c * S + r] = getp(R.Sr[j], c);
            }
            gam[(size_t)k * S + r] = R.gamma[j];
        }
    }
    const int nk = min(CF::KPC, N - k0);
    for (int e = threadIdx.x; e < nk * 3 * SS; e += CF::NT) {
        const int gg = e / (3 * SS), w = e - gg * 3 * SS;
        const int blk = w / SS, x = w - blk * SS, j = x / S, i = x - j * S;
        Pg[(size_t)(k0 + gg) * 3 * SS + w] = smem[CF::oP + (gg + 1) * CF::PK + i * CF::PR + prow_col<CF>(blk) + j];
    }
}

// load from the reference layout (load_block_bd semantics); the never-written corner blocks
// S[0][0], S[N-1][2], Pinv[0][0], Pinv[N-1][2] are NOT read.
template <class CF>
__device__ __forceinline__ void load_rows(typename CF::T* smem, RowsOf<CF>& R, const typename CF::T* __restrict__ Sg,
                                          const typename CF::T* __restrict__ Pg,
                                          const typename CF::T* __restrict__ gam, int N,
                                          bool glob_l = false, bool glob_r = false) {
    using T = typename CF::T;
    using T2 = typename CF::T2;
    // glob_l / glob_r: this cluster's first / last knot has a neighbour on ANOTHER rank (knot-range
    // split), so its left / right blocks are real data, not the never-written corner blocks
    constexpr int S = 14, SS = S * S;
    const int g = threadIdx.x / CF::LPK, t = threadIdx.x % CF::LPK;
    const int k0 = cluster_rank<CF::CL>() * CF::KPC;
    const int k = k0 + g;
    const RowMap<CF::LPK, CF::RPT, 14> rm(t);
    const T* s0 = Sg + (size_t)k * 3 * SS;
#pragma unroll
    for (int j = 0; j < CF::RPT; j++) {
        const bool on = (k < N) && rm.ok[j];
        const bool hl = on && (k > 0 || glob_l), hr = on && (k + 1 < N || glob_r);
        const int r = rm.row[j];
#pragma unroll
        for (int c = 0; c < S; c++) {
            setp(R.Sl[j], c, hl ? __ldg(s0 + 0 * SS + c * S + r) : T(0));
            setp(R.Sd[j], c, on ? __ldg(s0 + 1 * SS + c * S + r) : T(0));
            setp(R.Sr[j], c, hr ? __ldg(s0 + 2 * SS + c * S + r) : T(0));
        }
        R.gamma[j] = on ? __ldg(gam + (size_t)k * S + r) : T(0);
    }
    {
        constexpr int TOT2 = CF::KPC * 3 * SS / 2, ITERS = (TOT2 + CF::NT - 1) / CF::NT;
        const T2* src = reinterpret_cast<const T2*>(Pg + (size_t)k0 * 3 * SS);
        constexpr int UN = 8;     // loads first, stores after: UN loads in flight per lane
        for (int it0 = 0; it0 < ITERS; it0 += UN) {
            T2 v[UN];
#pragma unroll
            for (int u = 0; u < UN; u++) {
                const int e2 = threadIdx.x + (it0 + u) * CF::NT;
                const int e = 2 * e2;
                const int gg = e / (3 * SS), blk = (e - gg * 3 * SS) / SS;
                const int kk = k0 + gg;
                const bool corner = (kk == 0 && blk == 0 && !glob_l) || (kk == N - 1 && blk == 2 && !glob_r);
                v[u] = f2(T(0), T(0));
                if (it0 + u < ITERS && e2 < TOT2 && kk < N && !corner) v[u] = __ldg(src + e2);
            }
#pragma unroll
            for (int u = 0; u < UN; u++) {
                const int e2 = threadIdx.x + (it0 + u) * CF::NT;
                if (it0 + u < ITERS && e2 < TOT2) {
                    const int e = 2 * e2;
                    const int gg = e / (3 * SS), w = e - gg * 3 * SS;
                    const int blk = w / SS, x = w - blk * SS, j = x / S, i = x - j * S;
                    T* d = smem + CF::oP + (gg + 1) * CF::PK + i * CF::PR + prow_col<CF>(blk) + j;
                    d[0] = v[u].x; d[CF::PR] = v[u].y;
                }
            }
        }
    }
    for (int e = threadIdx.x; e < CF::KPC * S; e += CF::NT) {      // row padding (never used in sums)
        const int gg = e / S, i = e - gg * S;
        smem[CF::oP + (gg + 1) * CF::PK + i * CF::PR + 42] = T(0);
        smem[CF::oP + (gg + 1) * CF::PK + i * CF::PR + 43] = T(0);
    }
    __syncthreads();
}

template <class CF>
__global__ void __launch_bounds__(CF::NT, CF::MIN_CTAS) schur_kernel(BatchArgsT<typename CF::T> a) {
    extern __shared__ float4 smem4[];
    using T = typename CF::T;
    T* smem = reinterpret_cast<T*>(smem4);
    const int b = blockIdx.x / CF::CL;
    KktPtrs<T> P{a.G + b * a.sG, a.C + b * a.sC, a.g + b * a.sg, a.c + b * a.sc};
    RowsOf<CF> R;
    schur_rows<CF>(smem, P, a.N, a.rho, R);
    store_rows<CF>(smem, R, a.S + b * a.sS, a.Pinv + b * a.sS, a.gamma + b * a.sv, a.N);
}

template <class CF>
__global__ void __launch_bounds__(CF::NT, CF::MIN_CTAS) pcg_kernel(BatchArgsT<typename CF::T> a) {
    extern __shared__ float4 smem4[];
    using T = typename CF::T;
    T* smem = reinterpret_cast<T*>(smem4);
    const int b = blockIdx.x / CF::CL;
    RowsOf<CF> R;
    load_rows<CF>(smem, R, a.S + b * a.sS, a.Pinv + b * a.sS, a.gamma + b * a.sv, a.N);
    pcg_rows<CF>(smem, R, a.lambda + b * a.sv, a.N, a.max_iter, a.exit_tol, a.iters + b,
                 a.exits + b);
}

// The inputs (G, C, g, c, lambda0) of instance `b` -> L2 (prefetch.global.L2, one 128 B line per request).  With one CTA
// per SM nothing overlaps an instance's input staging (an HBM round trip per batch of loads: 10 % of the fused kernel's
// stall samples sat on it) with another instance's arithmetic; the 126 MB L2 holds the inputs of every instance that is
// about to start (148 x 215 KB = 32 MB), so the CTA that is iterating asks for the inputs of the instance that will get
// its SM next.  A hint only: no ordering, no correctness role.
template <class CF>
__device__ __forceinline__ void prefetch_inputs_l2(const BatchArgsT<typename CF::T>& a, long long b) {
    using T = typename CF::T;
    if (b >= a.batch) return;
    auto pf = [&](const T* base, size_t elems) {
        const char* p = reinterpret_cast<const char*>(base);
        const size_t bytes = elems * sizeof(T);
        for (size_t off = (size_t)threadIdx.x * 128; off < bytes; off += (size_t)CF::NT * 128)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(p + off));
    };
    pf(a.G + b * a.sG, a.sG); pf(a.C + b * a.sC, a.sC); pf(a.g + b * a.sg, a.sg); pf(a.c + b * a.sc, a.sc);
    pf(a.lambda + b * a.sv, a.sv);
}

template <class CF, int VAR = MPCB200_PCG_VARIANT>
__global__ void __launch_bounds__(CF::NT, CF::MIN_CTAS) fused_kernel(BatchArgsT<typename CF::T> a) {
    extern __shared__ float4 smem4[];
    using T = typename CF::T;
    T* smem = reinterpret_cast<T*>(smem4);
    const int b = blockIdx.x / CF::CL;
    KktPtrs<T> P{a.G + b * a.sG, a.C + b * a.sC, a.g + b * a.sg, a.c + b * a.sc};
    RowsOf<CF> R;
    schur_rows<CF>(smem, P, a.N, a.rho_dev ? __ldg(a.rho_dev + b * a.rho_stride) : a.rho, R);
    if (CF::CL == 1 && a.pf_dist > 0) prefetch_inputs_l2<CF>(a, (long long)b + a.pf_dist);
    pcg_rows<CF, false, VAR>(smem, R, a.lambda + b * a.sv, a.N, a.max_iter, a.exit_tol, a.iters + b,
                             a.exits + b);
    if (a.dz) {   // compute_dz epilogue: lambda and G^-1 were written by this cluster, read back through L2
        if (CF::CL == 1) __syncthreads();   // CL > 1: pcg_rows ended on a cluster barrier (release/acquire)
        dz_rows(P.G, P.C, P.g, a.lambda + b * a.sv, a.dz + b * a.sdz, a.N, (int)cluster_rank<CF::CL>() * CF::KPC, CF::KPC);
    }
}

// compute_dz alone (include/common/dz.cuh:124-133): 32 knots per 256-thread CTA over the whole batch
template <class T>
__global__ void __launch_bounds__(256) dz_kernel(BatchArgsT<T> a) {
    const int t = threadIdx.x & 7, lane = threadIdx.x & 31, base = lane & ~7;
    const unsigned mask = 0xFFu << base;
    const size_t total = (size_t)a.batch * a.N;
    for (size_t gk = (size_t)blockIdx.x * 32 + (threadIdx.x >> 3); gk < total; gk += (size_t)gridDim.x * 32) {
        const size_t b = gk / a.N;
        const int k = (int)(gk - b * a.N);
        dz_knot(a.G + b * a.sG, a.C + b * a.sC, a.g + b * a.sg, a.lambda + b * a.sv, a.dz + b * a.sdz, a.N, k, t, mask, base);
    }
}

// ---- knot-range split across GPUs (BASELINE.json configs[3]) -----------------------------------
// One cluster per RANK.  A real multi-GPU run launches one cluster per process/GPU; the
// single-GPU emulation used by the tests launches `ranks_here` clusters in ONE (cooperative)
// launch over all ranks' data -- the same code path with every mailbox local.
struct SplitArgs {
    int n_local;                 // knots per rank (== KPC * CL)
    unsigned grank_first, world, epoch;
    unsigned parity;             // half of every mailbox used by this solve (alternates between consecutive solves)
    unsigned long long timeout_ns;
    unsigned* diag;              // host-visible {flag, rank, expected tag, seen tag} of a timed-out poll (may be null)
    uint2* mail[SPLIT_MAXW];     // mailbox of every rank (local or peer-mapped)
    const float* S; const float* Pinv; const float* gamma; float* lambda;   // first local rank's slices
    unsigned max_iter; float exit_tol;
    unsigned* iters; uint8_t* exits;   // one entry per local rank
};

template <class CF>
__global__ void __launch_bounds__(CF::NT, CF::MIN_CTAS) pcg_split_kernel(SplitArgs a) {
    extern __shared__ float4 smem4[];
    float* smem = reinterpret_cast<float*>(smem4);
    const int c = blockIdx.x / CF::CL;                  // local rank index within this launch
    SplitCtx X{a.grank_first + (unsigned)c, a.world, a.epoch, a.mail, a.parity, a.timeout_ns, a.diag};
    const size_t ks = (size_t)c * a.n_local;
    RowsOf<CF> R;
    load_rows<CF>(smem, R, a.S + ks * 3 * 196, a.Pinv + ks * 3 * 196, a.gamma + ks * 14, a.n_local,
                  X.grank > 0, X.grank + 1 < X.world);
    pcg_rows<CF, true, 0>(smem, R, a.lambda + ks * 14, a.n_local, a.max_iter, a.exit_tol, a.iters + c,
                       a.exits + c, &X);
}

}  // namespace mpcb200
