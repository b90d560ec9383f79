// refstyle_pcg.cuh -- "reference-style" block-tridiagonal PCG: OUR restatement of the
// structure of the reference's (absent) GBD-PCG kernel as described in SURVEY 0.2:
//   cooperative launch, one CTA per knot, 4 grid-wide barriers per iteration, halo vectors and
//   per-knot partial sums round-tripping through global memory.
// It is NOT the reference's code (that source is not in the checkout) and must never be
// reported as "the reference".  It exists (a) as the comparison kernel the cluster/DSMEM
// kernels are measured against on the same box and (b) as the body of the header shim's
// pcg<T,STATE_SIZE,KNOT_POINTS> (mpcgpu_b200/shim/gpu_pcg.cuh), whose launch geometry
// (grid = knot_points, block = PCG_NUM_THREADS, cooperative) is fixed by the caller at
// include/pcg/sqp.cuh:275.
//
// Self-contained (no other project header) so the shim can include it on its own.
#pragma once
#include <cooperative_groups.h>
#include <cstdint>

namespace mpcb200_refstyle {

namespace cgrps = cooperative_groups;

template <typename T>
__host__ __device__ constexpr size_t smem_elems(uint32_t s) {
    // S row block + Pinv row block + gamma, lambda, r, p, ups, rt + halo vector + scratch
    return (size_t)6 * s * s + (size_t)6 * s + (size_t)3 * s + 64;
}

template <typename T>
__device__ __forceinline__ T warp_sum_t(T v) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// sum of n per-knot partials in global memory; every CTA does the same reduction in the
// same order => identical alpha / beta everywhere.  Result broadcast through s_scal[0].
template <typename T>
__device__ __forceinline__ T reduce_partials(const volatile T* d_part, uint32_t n, T* s_scal) {
    if (threadIdx.x < 32) {
        T acc = 0;
        for (uint32_t q = threadIdx.x; q < n; q += 32) acc += d_part[q];
        acc = warp_sum_t(acc);
        if (threadIdx.x == 0) s_scal[0] = acc;
    }
    __syncthreads();
    T out = s_scal[0];
    __syncthreads();
    return out;
}

// y_b = M_b [x_{b-1}; x_b; x_{b+1}]  (row block of M in shared memory, x halo in s_vec)
template <typename T>
__device__ __forceinline__ void bdmv_row(uint32_t s, const T* s_M, const T* s_vec, T* s_out) {
    for (uint32_t row = threadIdx.x; row < s; row += blockDim.x) {
        T acc = 0;
        for (uint32_t col = 0; col < 3 * s; col++) {
            const uint32_t blk = col / s, c = col - blk * s;
            acc += s_M[blk * s * s + c * s + row] * s_vec[col];
        }
        s_out[row] = acc;
    }
}

template <typename T>
__device__ __forceinline__ void load_halo(uint32_t s, uint32_t N, uint32_t b, const T* d_vec,
                                          T* s_vec) {
    for (uint32_t q = threadIdx.x; q < 3 * s; q += blockDim.x) {
        const int64_t src = (int64_t)b * s + q - s;
        s_vec[q] = (src >= 0 && src < (int64_t)N * s) ? __ldcg(d_vec + src) : (T)0;
    }
}

template <typename T>
__device__ __forceinline__ T block_dot(uint32_t s, const T* a, const T* b2) {
    // s <= 32: first warp only; result valid in thread 0
    T v = 0;
    if (threadIdx.x < 32) {
        for (uint32_t q = threadIdx.x; q < s; q += 32) v += a[q] * b2[q];
        v = warp_sum_t(v);
    }
    return v;
}

template <typename T>
__device__ void pcg_body(uint32_t s, uint32_t N, const T* d_S, const T* d_Pinv, const T* d_gamma,
                         T* d_lambda, T* d_r, T* d_p, T* d_v_temp, T* d_eta_new_temp,
                         uint32_t* d_iters, bool* d_max_iter_exit, uint32_t max_iter, T exit_tol,
                         T* smem) {
    cgrps::grid_group grid = cgrps::this_grid();
    const uint32_t b = blockIdx.x, ss = s * s;
    T* s_S = smem;
    T* s_P = s_S + 3 * ss;
    T* s_gamma = s_P + 3 * ss;
    T* s_lambda = s_gamma + s;
    T* s_r = s_lambda + s;
    T* s_p = s_r + s;
    T* s_ups = s_p + s;
    T* s_rt = s_ups + s;
    T* s_vec = s_rt + s;
    T* s_scal = s_vec + 3 * s;

    for (uint32_t q = threadIdx.x; q < 3 * ss; q += blockDim.x) {
        const uint32_t blk = q / ss;
        const bool corner = (b == 0 && blk == 0) || (b == N - 1 && blk == 2);   // never written
        s_S[q] = corner ? (T)0 : d_S[(size_t)b * 3 * ss + q];
        s_P[q] = corner ? (T)0 : d_Pinv[(size_t)b * 3 * ss + q];
    }
    for (uint32_t q = threadIdx.x; q < s; q += blockDim.x) {
        s_gamma[q] = d_gamma[(size_t)b * s + q];
        s_lambda[q] = d_lambda[(size_t)b * s + q];
    }
    load_halo(s, N, b, d_lambda, s_vec);
    __syncthreads();
    bdmv_row(s, s_S, s_vec, s_ups);
    __syncthreads();
    for (uint32_t q = threadIdx.x; q < s; q += blockDim.x) {
        s_r[q] = s_gamma[q] - s_ups[q];
        d_r[(size_t)b * s + q] = s_r[q];
    }
    grid.sync();
    load_halo(s, N, b, d_r, s_vec);
    __syncthreads();
    bdmv_row(s, s_P, s_vec, s_rt);
    __syncthreads();
    for (uint32_t q = threadIdx.x; q < s; q += blockDim.x) {
        s_p[q] = s_rt[q];
        d_p[(size_t)b * s + q] = s_p[q];
    }
    {
        T part = block_dot(s, s_r, s_rt);
        if (threadIdx.x == 0) d_eta_new_temp[b] = part;
    }
    grid.sync();
    T eta = reduce_partials(d_eta_new_temp, N, s_scal);

    uint32_t iter = 0;
    bool max_iter_exit = true;
    for (; iter < max_iter; iter++) {
        load_halo(s, N, b, d_p, s_vec);
        __syncthreads();
        bdmv_row(s, s_S, s_vec, s_ups);
        __syncthreads();
        {
            T part = block_dot(s, s_p, s_ups);
            if (threadIdx.x == 0) d_v_temp[b] = part;
        }
        grid.sync();                                                    // barrier 1
        const T v = reduce_partials(d_v_temp, N, s_scal);
        if (v == (T)0) { max_iter_exit = false; break; }
        const T alpha = eta / v;
        for (uint32_t q = threadIdx.x; q < s; q += blockDim.x) {
            s_lambda[q] += alpha * s_p[q];
            s_r[q] -= alpha * s_ups[q];
            d_r[(size_t)b * s + q] = s_r[q];
        }
        grid.sync();                                                    // barrier 2
        load_halo(s, N, b, d_r, s_vec);
        __syncthreads();
        bdmv_row(s, s_P, s_vec, s_rt);
        __syncthreads();
        {
            T part = block_dot(s, s_r, s_rt);
            if (threadIdx.x == 0) d_eta_new_temp[b] = part;
        }
        grid.sync();                                                    // barrier 3
        const T eta_new = reduce_partials(d_eta_new_temp, N, s_scal);
        const T aeta = eta_new < (T)0 ? -eta_new : eta_new;
        if (aeta < exit_tol) { iter++; max_iter_exit = false; break; }
        const T beta = eta_new / eta;
        eta = eta_new;
        for (uint32_t q = threadIdx.x; q < s; q += blockDim.x) {
            s_p[q] = s_rt[q] + beta * s_p[q];
            d_p[(size_t)b * s + q] = s_p[q];
        }
        grid.sync();                                                    // barrier 4
    }
    __syncthreads();
    for (uint32_t q = threadIdx.x; q < s; q += blockDim.x) d_lambda[(size_t)b * s + q] = s_lambda[q];
    if (b == 0 && threadIdx.x == 0) { d_iters[0] = iter; d_max_iter_exit[0] = max_iter_exit; }
}

}  // namespace mpcb200_refstyle
