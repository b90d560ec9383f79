// glass.cuh -- drop-in for the (absent) GLASS submodule: only the four block-level primitives
// the in-tree reference code calls.  Semantics pinned by the call sites:
//   copy(n, src, dst) / copy(n, alpha, src, dst)   include/common/kkt.cuh:109-117
//   gemm<T,TRANSPOSE_B>(m, n, k, alpha, A, B, C)   include/pcg/linsys_setup.cuh:398,406,447,473
//   reduce(n, x)                                   include/common/merit.cuh:74,88
// All are block-strided (every thread of the CTA calls them); none synchronises on exit --
// the callers place their own __syncthreads().
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace glass {

template <typename T>
__device__ void copy(uint32_t n, const T* src, T* dst) {
    for (uint32_t e = threadIdx.x; e < n; e += blockDim.x) dst[e] = src[e];
}

template <typename T>
__device__ void copy(uint32_t n, T alpha, const T* src, T* dst) {
    for (uint32_t e = threadIdx.x; e < n; e += blockDim.x) dst[e] = alpha * src[e];
}

// column-major: A is m x n, B is n x k (k x n when TRANSPOSE_B), C = alpha * A * op(B) is m x k
template <typename T, bool TRANSPOSE_B = false>
__device__ void gemm(uint32_t m, uint32_t n, uint32_t k, T alpha, const T* A, const T* B, T* C) {
    for (uint32_t e = threadIdx.x; e < m * k; e += blockDim.x) {
        const uint32_t col = e / m, row = e - col * m;
        T acc = static_cast<T>(0);
        for (uint32_t l = 0; l < n; l++)
            acc += A[row + l * m] * (TRANSPOSE_B ? B[col + l * k] : B[l + col * n]);
        C[e] = alpha * acc;
    }
}

// in-place sum of x[0..n), result in x[0]; any n; ends synchronised
template <typename T>
__device__ void reduce(uint32_t n, T* x) {
    __syncthreads();
    while (n > 1) {
        const uint32_t half = (n + 1) / 2;
        for (uint32_t e = threadIdx.x; e + half < n; e += blockDim.x) x[e] += x[e + half];
        __syncthreads();
        n = half;
    }
}

template <typename T>
__device__ T dot(uint32_t n, const T* a, const T* b, T* scratch) {
    for (uint32_t e = threadIdx.x; e < n; e += blockDim.x) scratch[e] = a[e] * b[e];
    reduce<T>(n, scratch);
    return scratch[0];
}

}  // namespace glass
