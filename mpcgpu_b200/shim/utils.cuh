// utils.cuh -- drop-in for GBD-PCG's utils.cuh: the block-tridiagonal buffer accessors the
// reference's form_schur kernels call.  Signatures pinned verbatim by the forward declarations
// at include/pcg/linsys_setup_no_precond.cuh:29-49; the buffer is [knot_points][3][s*s]
// column-major blocks (left, diag, right), include/pcg/sqp.cuh:111,130.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

// dst[(block_row*3 + block_col)*s*s + ...] = multiplier * src   (block-strided over the CTA)
template <typename T>
__device__ void store_block_bd(uint32_t state_size, uint32_t knot_points, const T* src, T* dst,
                               unsigned block_col, unsigned block_row, int8_t multiplier = 1) {
    (void)knot_points;
    const uint32_t ss = state_size * state_size;
    T* out = dst + ((size_t)block_row * 3 + block_col) * ss;
    const T m = static_cast<T>(multiplier);
    for (uint32_t e = threadIdx.x; e < ss; e += blockDim.x) out[e] = m * src[e];
}

// dst = block (block_row, block_col) of src, optionally transposed
template <typename T>
__device__ void load_block_bd(uint32_t state_size, uint32_t knot_points, const T* src, T* dst,
                              unsigned block_col, unsigned block_row, bool transpose = false) {
    (void)knot_points;
    const uint32_t ss = state_size * state_size;
    const T* in = src + ((size_t)block_row * 3 + block_col) * ss;
    for (uint32_t e = threadIdx.x; e < ss; e += blockDim.x) {
        const uint32_t col = e / state_size, row = e - col * state_size;
        dst[transpose ? row * state_size + col : e] = in[e];
    }
}

// block-strided copy, DESTINATION first (include/qdldl/linsys_setup.cuh:67-70)
template <typename T>
__device__ void gato_memcpy(T* dst, const T* src, unsigned n) {
    for (unsigned e = threadIdx.x; e < n; e += blockDim.x) dst[e] = src[e];
}
