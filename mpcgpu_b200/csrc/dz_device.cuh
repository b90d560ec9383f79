// dz_device.cuh -- primal step recovery from lambda (the consumer of the hot
// path's outputs): the math of compute_dz_kernel, include/common/dz.cuh:3-121
//     state rows   : dz_x[k] = Qi_k ( q_k - lambda_k - Abar_k^T lambda_{k+1} )
//     control rows : dz_u[k] = Ri_k ( r_k - Bbar_k^T lambda_{k+1} )            (k < N-1)
// with Qi, Ri the blocks of G^-1 that form_schur_system left in d_G.  Eight lanes per knot: lane t
// owns state rows t, t+8 and control row t; columns of Abar/Bbar are contiguous (float2 loads), rows
// of Qi/Ri are read coalesced across the eight lanes; the intermediate vector moves by shuffles.
#pragma once
#include "common.cuh"

namespace mpcb200 {

// lambda is read through L2 (__ldcg): in the fused kernel it was written moments ago by this or a
// neighbouring CTA of the cluster, and L1 may still hold the warm-start values.
__device__ __forceinline__ void dz_knot(const float* __restrict__ Ginv, const float* __restrict__ C,
                                        const float* __restrict__ g, const float* lam, float* __restrict__ dz,
                                        int N, int k, int t, unsigned mask, int base) {
    constexpr int S = 14, U = 7;
    const bool last = (k == N - 1);
    float ln[S];
    if (!last) {
        const float2* p = reinterpret_cast<const float2*>(lam + (size_t)(k + 1) * S);
#pragma unroll
        for (int j = 0; j < S / 2; j++) { float2 v = __ldcg(p + j); ln[2 * j] = v.x; ln[2 * j + 1] = v.y; }
    } else {
#pragma unroll
        for (int j = 0; j < S; j++) ln[j] = 0.f;
    }
    const float* gk = g + (size_t)k * (S + U);
    const float* Ck = C + (size_t)k * (S * S + S * U);   // never dereferenced when last
    float tv[2], uv = 0.f;
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const int i = t + 8 * h;
        float acc = 0.f;
        if (i < S && !last) {
            const float2* col = reinterpret_cast<const float2*>(Ck + i * S);
#pragma unroll
            for (int j = 0; j < S / 2; j++) { float2 a = __ldg(col + j); acc = fmaf(a.x, ln[2 * j], acc); acc = fmaf(a.y, ln[2 * j + 1], acc); }
        }
        tv[h] = (i < S) ? __ldg(gk + i) - (__ldcg(lam + (size_t)k * S + i) + acc) : 0.f;
    }
    if (t < U && !last) {
        const float2* col = reinterpret_cast<const float2*>(Ck + S * S + t * S);
        float acc = 0.f;
#pragma unroll
        for (int j = 0; j < S / 2; j++) { float2 a = __ldg(col + j); acc = fmaf(a.x, ln[2 * j], acc); acc = fmaf(a.y, ln[2 * j + 1], acc); }
        uv = __ldg(gk + S + t) - acc;
    }
    const float* Qi = Ginv + (size_t)k * (S * S + U * U);
    float x0 = 0.f, x1 = 0.f, u0 = 0.f;
#pragma unroll
    for (int j = 0; j < S; j++) {
        const float tj = __shfl_sync(mask, j < 8 ? tv[0] : tv[1], base + (j & 7));
        x0 = fmaf(__ldcg(Qi + j * S + t), tj, x0);
        if (t < S - 8) x1 = fmaf(__ldcg(Qi + j * S + t + 8), tj, x1);
    }
#pragma unroll
    for (int j = 0; j < U; j++) {
        const float uj = __shfl_sync(mask, uv, base + j);
        if (t < U && !last) u0 = fmaf(__ldcg(Qi + S * S + j * U + t), uj, u0);
    }
    float* out = dz + (size_t)k * (S + U);
    out[t] = x0;
    if (t < S - 8) out[t + 8] = x1;
    if (t < U && !last) out[S + t] = u0;
}

// knots [k0, k0+nk) of one instance by the threads of one CTA (blockDim.x a multiple of 32)
__device__ __forceinline__ void dz_rows(const float* Ginv, const float* C, const float* g, const float* lam,
                                        float* dz, int N, int k0, int nk) {
    const int t = threadIdx.x & 7, lane = threadIdx.x & 31, base = lane & ~7;
    const unsigned mask = 0xFFu << base;
    for (int kk = threadIdx.x >> 3; kk < nk; kk += blockDim.x >> 3)
        if (k0 + kk < N) dz_knot(Ginv, C, g, lam, dz, N, k0 + kk, t, mask, base);
}

}  // namespace mpcb200
