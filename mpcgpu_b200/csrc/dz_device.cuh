// dz_device.cuh -- primal step recovery from lambda (the consumer of the hot
// path's outputs): the math of compute_dz_kernel, include/common/dz.cuh:3-121
//     state rows   : dz_x[k] = Qi_k ( q_k - lambda_k - Abar_k^T lambda_{k+1} )
//     control rows : dz_u[k] = Ri_k ( r_k - Bbar_k^T lambda_{k+1} )            (k < N-1)
// with Qi, Ri the blocks of G^-1 that form_schur_system left in d_G.  Eight lanes per knot: lane t
// owns state rows t, t+8 and control row t; columns of Abar/Bbar are contiguous (float2 loads), rows
// of Qi/Ri are read coalesced across the eight lanes; the intermediate vector moves by shuffles.
#pragma once
#include "kkt_device.cuh"

namespace mpcb200 {

// lambda is read through L2 (__ldcg): in the fused kernel it was written moments ago by this or a
// neighbouring CTA of the cluster, and L1 may still hold the warm-start values.
template <class T>
__device__ __forceinline__ void dz_knot(const T* __restrict__ Ginv, const T* __restrict__ C,
                                        const T* __restrict__ g, const T* lam, T* __restrict__ dz,
                                        int N, int k, int t, unsigned mask, int base) {
    using T2 = typename Num<T>::T2;
    constexpr int S = 14, U = 7;
    const bool last = (k == N - 1);
    T ln[S];
    if (!last) {
        const T2* p = reinterpret_cast<const T2*>(lam + (size_t)(k + 1) * S);
#pragma unroll
        for (int j = 0; j < S / 2; j++) { T2 v = __ldcg(p + j); ln[2 * j] = v.x; ln[2 * j + 1] = v.y; }
    } else {
#pragma unroll
        for (int j = 0; j < S; j++) ln[j] = T(0);
    }
    const T* gk = g + (size_t)k * (S + U);
    const T* Ck = C + (size_t)k * (S * S + S * U);   // never dereferenced when last
    T tv[2], uv = T(0);
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const int i = t + 8 * h;
        T acc = T(0);
        if (i < S && !last) {
            const T2* col = reinterpret_cast<const T2*>(Ck + i * S);
#pragma unroll
            for (int j = 0; j < S / 2; j++) { T2 a = __ldg(col + j); acc = fm(a.x, ln[2 * j], acc); acc = fm(a.y, ln[2 * j + 1], acc); }
        }
        tv[h] = (i < S) ? __ldg(gk + i) - (__ldcg(lam + (size_t)k * S + i) + acc) : T(0);
    }
    if (t < U && !last) {
        const T2* col = reinterpret_cast<const T2*>(Ck + S * S + t * S);
        T acc = T(0);
#pragma unroll
        for (int j = 0; j < S / 2; j++) { T2 a = __ldg(col + j); acc = fm(a.x, ln[2 * j], acc); acc = fm(a.y, ln[2 * j + 1], acc); }
        uv = __ldg(gk + S + t) - acc;
    }
    const T* Qi = Ginv + (size_t)k * (S * S + U * U);
    T x0 = T(0), x1 = T(0), u0 = T(0);
#pragma unroll
    for (int j = 0; j < S; j++) {
        const T tj = __shfl_sync(mask, j < 8 ? tv[0] : tv[1], base + (j & 7));
        x0 = fm(__ldcg(Qi + j * S + t), tj, x0);
        if (t < S - 8) x1 = fm(__ldcg(Qi + j * S + t + 8), tj, x1);
    }
#pragma unroll
    for (int j = 0; j < U; j++) {
        const T uj = __shfl_sync(mask, uv, base + j);
        if (t < U && !last) u0 = fm(__ldcg(Qi + S * S + j * U + t), uj, u0);
    }
    T* out = dz + (size_t)k * (S + U);
    out[t] = x0;
    if (t < S - 8) out[t + 8] = x1;
    if (t < U && !last) out[S + t] = u0;
}

// knots [k0, k0+nk) of one instance by the threads of one CTA (blockDim.x a multiple of 32)
template <class T>
__device__ __forceinline__ void dz_rows(const T* Ginv, const T* C, const T* g, const T* lam,
                                        T* dz, int N, int k0, int nk) {
    const int t = threadIdx.x & 7, lane = threadIdx.x & 31, base = lane & ~7;
    const unsigned mask = 0xFFu << base;
    for (int kk = threadIdx.x >> 3; kk < nk; kk += blockDim.x >> 3)
        if (k0 + kk < N) dz_knot(Ginv, C, g, lam, dz, N, k0 + kk, t, mask, base);
}

}  // namespace mpcb200
