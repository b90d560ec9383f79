// b200/kkt_batched.cuh -- batched producer of the hot path's inputs:
// the per-knot KKT blocks G = [Q_k | R_k], C = [-A_k | -B_k], g = [q_k | r_k], c of MANY independent
// MPC instances (BASELINE.json configs[4]: "perturbed start states") in ONE launch, written in the
// batched layout of include/mpcgpu_b200.h (instance b at base + b * <single-instance size>), so
// that   generate_kkt_submatrices_batched -> mpcgpu_b200_kkt_solve_dz_f32(batch)   is an
// end-to-end linearise-and-solve step with nothing but the trajectories resident beforehand.
//
// What it computes per (instance, knot) is exactly what one block of the reference's
// generate_kkt_submatrices (include/common/kkt.cuh:22-163) computes, with the reference's own
// dynamics and cost device functions called from where they are (GRiD is kept as is):
//   integratorAndGradient<T, INTEGRATOR_TYPE, ANGLE_WRAP, true>   include/common/integrator.cuh:133-157
//   gato_plant::trackingCostGradientAndHessian[_lastblock]<T>     include/dynamics/rbd_plant.cuh
// Those functions cooperate over threadIdx only, so a block here produces the bits the reference's
// block produces for the same knot (tools/kkt_batched_check.cu asserts it).  Different from the
// reference kernel: a 2-D grid (knot, instance) instead of a grid-stride loop over one instance; one
// code path for the inner and the last knot pair (the last one additionally emits Q_N, q_N); the
// stores are plain strided loops (no glass:: dependency).
//
// Inputs per instance b (strides in elements):
//   d_xu    + b * xu_stride    states and controls [x_0 u_0 x_1 u_1 ... x_{N-1}], (s+c)*N - c values
//   d_xs    + b * s            the measured start state (c_0 = x_0 - xs)
//   d_eePos + b * ee_stride    end-effector goal, 6 values per knot (ee_stride = 0: one shared goal)
#pragma once
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>
#include <cooperative_groups.h>
#ifndef MPCGPU_B200_SLOT_COMMON_INCLUDED   // common/kkt.cuh carries no include guard
#include "common/kkt.cuh"                  // get_kkt_smem_size, integrator.cuh, rbd_plant.cuh
#endif

namespace b200slot {

// threads per block of the batched dynamics-facing kernels.  The reference's device functions stride over
// blockDim.x, so any multiple of 32 computes the same bits; with thousands of blocks in flight smaller blocks
// than the reference's 128 (KKT_THREADS, MERIT_THREADS) are faster -- measured on B200, 4096 x N=64:
// KKT producer 16.6 ms (128) / 15.6 (64) / 18.7 (32); line-search merits: SQP iteration 49 ms (128) / 43 (64) /
// 37 (32).  Environment variables B200SLOT_KKT_THREADS / B200SLOT_MERIT_THREADS override.
inline unsigned batched_threads(const char* env, unsigned dflt) {
    const char* e = getenv(env);
    const int v = e ? atoi(e) : 0;
    return (v >= 32 && v <= 1024 && v % 32 == 0) ? (unsigned)v : dflt;
}

template <typename T>
__device__ __forceinline__ void block_store(uint32_t n, const T* __restrict__ src, T* __restrict__ dst, T scale) {
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) dst[i] = scale * src[i];
}

template <typename T, unsigned INTEGRATOR_TYPE = 0, bool ANGLE_WRAP = false>
__global__ void generate_kkt_submatrices_batched(uint32_t state_size, uint32_t control_size, uint32_t knot_points,
                                                 T* d_G_dense, T* d_C_dense, T* d_g, T* d_c, void* d_dynMem_const,
                                                 T timestep, const T* d_eePos_traj, size_t ee_stride, const T* d_xs,
                                                 const T* d_xu, size_t xu_stride) {
    namespace cgrps = cooperative_groups;
    const cgrps::thread_block block = cgrps::this_thread_block();
    const uint32_t s = state_size, c = control_size, N = knot_points;
    const uint32_t ss = s * s, sc = s * c, cc = c * c, spc = s + c;
    const uint32_t k = blockIdx.x;            // knot pair (k, k+1), k < N-1
    const size_t b = blockIdx.y;              // instance
    if (k + 1 >= N) return;
    const bool last = (k == N - 2);

    // this instance's slices (sizes of include/pcg/sqp.cuh:54-58)
    T* G = d_G_dense + b * ((size_t)(ss + cc) * N - cc);
    T* C = d_C_dense + b * ((size_t)(ss + sc) * (N - 1));
    T* g = d_g + b * ((size_t)spc * N - c);
    T* cv = d_c + b * ((size_t)s * N);
    const T* xu = d_xu + b * xu_stride;
    const T* ee = d_eePos_traj + b * ee_stride;
    const T* xs = d_xs + b * s;

    // shared-memory carve-up: the reference's own (its device functions size their scratch from it,
    // get_kkt_smem_size, include/common/kkt.cuh:5-19)
    extern __shared__ unsigned char kkt_batched_smem[];
    T* s_xux = reinterpret_cast<T*>(kkt_batched_smem);
    T* s_ee = s_xux + 2 * s + c;
    T* s_Qk = s_ee + 6;
    T* s_Rk = s_Qk + ss;
    T* s_qk = s_Rk + cc;
    T* s_rk = s_qk + s;
    T* s_Ak = s_rk + c;
    T* s_Bk = s_Ak + ss;
    T* s_Qn = s_Bk + sc;                               // Q_N, q_N: last knot pair only
    T* s_qn = s_Qn + ss;
    T* s_err = last ? s_qn + s : s_Bk + sc;            // integrator defect x_{k+1} - f(x_k, u_k)
    T* s_tmp = s_err + s;

    for (uint32_t i = threadIdx.x; i < 2 * s + c; i += blockDim.x) s_xux[i] = xu[(size_t)k * spc + i];
    for (uint32_t i = threadIdx.x; i < 12; i += blockDim.x) s_ee[i] = ee[(size_t)k * 6 + i];
    __syncthreads();

    integratorAndGradient<T, INTEGRATOR_TYPE, ANGLE_WRAP, true>(s, c, s_xux, s_Ak, s_Bk, s_err, s_tmp, d_dynMem_const,
                                                                timestep, block);
    __syncthreads();
    if (last)
        gato_plant::trackingCostGradientAndHessian_lastblock<T>(s, c, s_xux, s_ee, s_Qk, s_qk, s_Rk, s_rk, s_Qn, s_qn,
                                                                s_tmp, d_dynMem_const);
    else
        gato_plant::trackingCostGradientAndHessian<T>(s, c, s_xux, s_ee, s_Qk, s_qk, s_Rk, s_rk, s_tmp, d_dynMem_const);
    __syncthreads();

    block_store<T>(ss, s_Qk, G + (size_t)(ss + cc) * k, T(1));
    block_store<T>(cc, s_Rk, G + (size_t)(ss + cc) * k + ss, T(1));
    block_store<T>(s, s_qk, g + (size_t)spc * k, T(1));
    block_store<T>(c, s_rk, g + (size_t)spc * k + s, T(1));
    block_store<T>(ss, s_Ak, C + (size_t)(ss + sc) * k, T(-1));          // Abar = -A, Bbar = -B
    block_store<T>(sc, s_Bk, C + (size_t)(ss + sc) * k + ss, T(-1));
    block_store<T>(s, s_err, cv + (size_t)s * (k + 1), T(1));
    if (last) {
        block_store<T>(ss, s_Qn, G + (size_t)(ss + cc) * (k + 1), T(1));
        block_store<T>(s, s_qn, g + (size_t)spc * (k + 1), T(1));
        for (uint32_t i = threadIdx.x; i < s; i += blockDim.x) cv[i] = xu[i] - xs[i];   // c_0 (kkt.cuh:103-105)
    }
}

// launch: grid (N-1, batch), the reference's block size and shared memory (include/pcg/sqp.cuh:66,213)
template <typename T>
inline cudaError_t launch_kkt_batched(uint32_t state_size, uint32_t control_size, uint32_t knot_points, uint32_t batch,
                                      T* d_G_dense, T* d_C_dense, T* d_g, T* d_c, void* d_dynMem_const, T timestep,
                                      const T* d_eePos_traj, size_t ee_stride, const T* d_xs, const T* d_xu,
                                      size_t xu_stride, cudaStream_t stream) {
    if (knot_points < 2 || batch == 0 || batch > 65535) return cudaErrorInvalidValue;
    const size_t smem = 2 * get_kkt_smem_size<T>(state_size, control_size);
    auto kern = generate_kkt_submatrices_batched<T>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    kern<<<dim3(knot_points - 1, batch), batched_threads("B200SLOT_KKT_THREADS", 64), smem, stream>>>(state_size, control_size, knot_points, d_G_dense,
                                                                    d_C_dense, d_g, d_c, d_dynMem_const, timestep,
                                                                    d_eePos_traj, ee_stride, d_xs, d_xu, xu_stride);
    return cudaGetLastError();
}

}  // namespace b200slot
