// kernels.cuh -- __global__ entry points built from kkt_device.cuh.
//
// Grid: batch * CL CTAs, one thread-block cluster of CL CTAs per KKT instance, KPC knots per
// CTA, 16 lanes per knot.  Instance b's buffers start at base + b * <elements per instance>
// (the reference's single-instance buffers, concatenated).
#pragma once
#include "kkt_device.cuh"

namespace mpcb200 {

struct BatchArgs {
    int N;
    int batch;
    // per-instance element strides
    size_t sG, sC, sg, sc, sS, sv;
    float* G; const float* C; const float* g; const float* c;
    float rho;
    float* S; float* Pinv; float* gamma;   // two-stage outputs / inputs
    float* lambda;
    unsigned max_iter; float exit_tol;
    unsigned* iters; uint8_t* exits;
};

// store my rows to the reference [N][3][S*S] column-major layout (store_block_bd semantics,
// include/pcg/linsys_setup_no_precond.cuh:29-38); corner blocks are written as zeros.
template <class CF>
__device__ __forceinline__ void store_rows(const Rows<CF::S>& R, float* Sg, float* Pg, float* gam,
                                           int N) {
    constexpr int S = CF::S, SS = S * S;
    const int g = threadIdx.x / LPK, i = threadIdx.x % LPK;
    const int k = cluster_rank<CF::CL>() * CF::KPC + g;
    if (k >= N || i >= S) return;
    float* s0 = Sg + (size_t)k * 3 * SS + i;
    float* p0 = Pg + (size_t)k * 3 * SS + i;
#pragma unroll
    for (int j = 0; j < S; j++) {
        s0[0 * SS + j * S] = R.Sl[j]; s0[1 * SS + j * S] = R.Sd[j]; s0[2 * SS + j * S] = R.Sr[j];
        p0[0 * SS + j * S] = R.Pl[j]; p0[1 * SS + j * S] = R.Pd[j]; p0[2 * SS + j * S] = R.Pr[j];
    }
    gam[(size_t)k * S + i] = R.gamma;
}

// load my rows from the reference layout (load_block_bd semantics); the never-written corner
// blocks S[0][0], S[N-1][2], Pinv[0][0], Pinv[N-1][2] are NOT read.
template <class CF>
__device__ __forceinline__ void load_rows(Rows<CF::S>& R, const float* __restrict__ Sg,
                                          const float* __restrict__ Pg,
                                          const float* __restrict__ gam, int N) {
    constexpr int S = CF::S, SS = S * S;
    const int g = threadIdx.x / LPK, i = threadIdx.x % LPK;
    const int k = cluster_rank<CF::CL>() * CF::KPC + g;
    const bool on = (k < N) && (i < S);
    const bool hl = on && k > 0, hr = on && (k + 1 < N);
    const float* s0 = Sg + (size_t)k * 3 * SS + i;
    const float* p0 = Pg + (size_t)k * 3 * SS + i;
#pragma unroll
    for (int j = 0; j < S; j++) {
        R.Sl[j] = hl ? __ldg(s0 + 0 * SS + j * S) : 0.f;
        R.Sd[j] = on ? __ldg(s0 + 1 * SS + j * S) : 0.f;
        R.Sr[j] = hr ? __ldg(s0 + 2 * SS + j * S) : 0.f;
        R.Pl[j] = hl ? __ldg(p0 + 0 * SS + j * S) : 0.f;
        R.Pd[j] = on ? __ldg(p0 + 1 * SS + j * S) : 0.f;
        R.Pr[j] = hr ? __ldg(p0 + 2 * SS + j * S) : 0.f;
    }
    R.gamma = on ? __ldg(gam + (size_t)k * S + i) : 0.f;
}

template <class CF>
__global__ void __launch_bounds__(CF::NT, 1) schur_kernel(BatchArgs a) {
    extern __shared__ float4 smem4[];
    float* smem = reinterpret_cast<float*>(smem4);
    const int b = blockIdx.x / CF::CL;
    KktPtrs P{a.G + b * a.sG, a.C + b * a.sC, a.g + b * a.sg, a.c + b * a.sc};
    Rows<CF::S> R;
    schur_rows<CF>(smem, P, a.N, a.rho, R);
    store_rows<CF>(R, a.S + b * a.sS, a.Pinv + b * a.sS, a.gamma + b * a.sv, a.N);
    if (CF::CL > 1) cluster_sync_all<CF::CL>();
}

template <class CF>
__global__ void __launch_bounds__(CF::NT, 1) pcg_kernel(BatchArgs a) {
    extern __shared__ float4 smem4[];
    float* smem = reinterpret_cast<float*>(smem4);
    const int b = blockIdx.x / CF::CL;
    Rows<CF::S> R;
    load_rows<CF>(R, a.S + b * a.sS, a.Pinv + b * a.sS, a.gamma + b * a.sv, a.N);
    pcg_rows<CF>(smem, R, a.lambda + b * a.sv, a.N, a.max_iter, a.exit_tol, a.iters + b,
                 a.exits + b);
}

// fused KKT solve: S and Pinv never leave the SM (Def. B traffic: read G,C,g,c,lambda0;
// write lambda, G^-1).
template <class CF>
__global__ void __launch_bounds__(CF::NT, 1) fused_kernel(BatchArgs a) {
    extern __shared__ float4 smem4[];
    float* smem = reinterpret_cast<float*>(smem4);
    const int b = blockIdx.x / CF::CL;
    KktPtrs P{a.G + b * a.sG, a.C + b * a.sC, a.g + b * a.sg, a.c + b * a.sc};
    Rows<CF::S> R;
    schur_rows<CF>(smem, P, a.N, a.rho, R);
    pcg_rows<CF>(smem, R, a.lambda + b * a.sv, a.N, a.max_iter, a.exit_tol, a.iters + b,
                 a.exits + b);
}

}  // namespace mpcb200
