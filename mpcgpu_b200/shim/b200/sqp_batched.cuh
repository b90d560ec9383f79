// b200/sqp_batched.cuh -- SQP iterations of MANY independent MPC instances with no host in the loop
// ("line-search merit batching (8 alphas x instances in one launch) and ... a
// fused update"; BASELINE.json configs[4]).  One iteration of include/pcg/sqp.cuh:203-420 for B
// instances is FOUR launches on one stream and no read-back:
//
//   1. b200slot::launch_kkt_batched                  linearise   (b200/kkt_batched.cuh)
//   2. mpcgpu_b200_kkt_solve_dz_rhovec_f32            Schur + PCG + compute_dz, rho per instance
//   3. b200slot::ls_merit_batched                    the 8 step lengths x N knots x B instances of the
//                                                    parallel line search in one launch (the reference
//                                                    launches 8 cooperative kernels on 8 streams per
//                                                    instance, include/pcg/sqp.cuh:327-346)
//   4. b200slot::select_step_batched                 per instance: sum the point merits, pick the step
//                                                    length, adapt rho / drho, xu += alpha dz
//                                                    (the host logic of include/pcg/sqp.cuh:352-413 and
//                                                    its cublasSaxpy, on the device)
//
// The per-point arithmetic is the reference's own: gato_plant::trackingcost and integratorError are
// called from where they are exactly as ls_gato_compute_merit / compute_merit call them
// (include/common/merit.cuh:17-94,96-141), the knot sum uses the same glass::reduce, and the decision
// rule is the reference's, so an instance follows the trajectory the reference's loop would give it
// (tools/sqp_batched_check.cu asserts xu, rho and merit bit for bit against that loop).
// Differences by construction: the initial merit is a fixed-order sum (the reference accumulates it
// with atomicAdd, order not defined); there is no wall-clock budget inside an iteration (the caller
// decides how many iterations to enqueue).
#pragma once
#include <cstdint>
#include <type_traits>
#include <vector>
#include <cuda_runtime.h>
#include <cooperative_groups.h>
#include "settings.cuh"
#ifndef MPCGPU_B200_SLOT_COMMON_INCLUDED
#include "common/kkt.cuh"
#endif
#include "merit.cuh"
#include "glass.cuh"
#include "b200/kkt_batched.cuh"
#include "mpcgpu_b200.h"

namespace b200slot {

constexpr uint32_t kLsAlphas = 8;     // step lengths 1, 1/2, ... 1/128 (include/pcg/sqp.cuh:63,410)

// Per-instance SQP state, all in device memory (arrays of `batch`).
template <typename T>
struct BatchedSqpState {
    T* rho;          // regularisation                      (include/pcg/sqp.cuh:97, mpcsim.cuh:256)
    T* drho;         // its growth factor                   (include/pcg/sqp.cuh:96)
    T* merit;        // merit of the current iterate
    T* step;         // step length taken by the last iteration (0: none, rho raised instead)
    uint32_t* flags; // bit 0: rho exceeded RHO_MAX and was reset.  The reference LEAVES its SQP loop at that point
                     // (include/pcg/sqp.cuh:379-383): from then on the instance is frozen here too -- select_step_batched
                     // takes no further step and changes neither rho nor the merit (the batched linearise / solve / line
                     // search kernels still compute for it: their results are ignored), until the caller clears the bit
                     // (the reference's next sqpSolvePcg call starts again from rho_reset).
};

// Point merits J_k + mu c_k of (knot, step length p, instance): slot p < 8 evaluates xu + alpha_p dz like
// ls_gato_compute_merit (alpha_p = -1 / 2^p, start-state defect on the last knot); slot 8 evaluates xu
// itself like compute_merit (no start-state term).  d_point: [batch][9][knot_points].
template <typename T>
__global__ void ls_merit_batched(uint32_t state_size, uint32_t control_size, uint32_t knot_points, const T* d_xs,
                                 const T* d_xu, size_t xu_stride, const T* d_eePos_traj, size_t ee_stride, T mu, T dt,
                                 void* d_dynMem_const, const T* d_dz, uint32_t first_slot, T* d_point) {
    grid::robotModel<T>* d_robotModel = (grid::robotModel<T>*)d_dynMem_const;
    const cooperative_groups::thread_block block = cooperative_groups::this_thread_block();
    const uint32_t knot = blockIdx.x, slot = first_slot + blockIdx.y;
    const size_t b = blockIdx.z;
    const uint32_t spc = state_size + control_size;
    const T* xu = d_xu + b * xu_stride;
    const T* dz = d_dz + b * xu_stride;
    const T* ee = d_eePos_traj + b * ee_stride;
    const T* xs = d_xs + b * state_size;
    const bool ls = slot < kLsAlphas;
    const T alpha = ls ? T(-1.0 / (1 << slot)) : T(0);

    extern __shared__ unsigned char ls_batched_smem[];
    T* s_xux_k = reinterpret_cast<T*>(ls_batched_smem);
    T* s_eePos_k_traj = s_xux_k + 2 * state_size + control_size;
    T* s_temp = s_eePos_k_traj + 6;

    for (int i = threadIdx.x; i < (int)(state_size + (knot < knot_points - 1) * spc); i += blockDim.x) {
        T x = xu[knot * spc + i];
        if (ls) x = x + alpha * dz[knot * spc + i];
        s_xux_k[i] = x;
        if (i < 6) s_eePos_k_traj[i] = ee[knot * 6 + i];
    }
    block.sync();
    const T Jk = gato_plant::trackingcost<T>(state_size, control_size, knot_points, s_xux_k, s_eePos_k_traj, s_temp, d_robotModel);
    block.sync();
    T ck = T(0);
    if (knot < knot_points - 1) {
        ck = integratorError<T>(state_size, s_xux_k, &s_xux_k[spc], s_temp, d_robotModel, dt, block);
    } else if (ls) {
        // |x_0 + alpha dz_0 - xs| summed (include/common/merit.cuh:68-76; every thread writes the same values there)
        for (uint32_t i = threadIdx.x; i < state_size; i += blockDim.x) s_temp[i] = abs((xu[i] + alpha * dz[i]) - xs[i]);
        block.sync();
        glass::reduce<T>(state_size, s_temp);
        block.sync();
        ck = s_temp[0];
    }
    block.sync();
    if (threadIdx.x == 0) d_point[(b * (kLsAlphas + 1) + slot) * knot_points + knot] = Jk + mu * ck;
}

// initial merit of every instance from slot 8 of d_point (after ls_merit_batched with first_slot = 8)
template <typename T>
__global__ void init_merit_batched(uint32_t knot_points, T* d_point, T* d_merit) {
    T* p = d_point + ((size_t)blockIdx.x * (kLsAlphas + 1) + kLsAlphas) * knot_points;
    glass::reduce<T>(knot_points, p);
    if (threadIdx.x == 0) d_merit[blockIdx.x] = p[0];
}

// One block per instance: merits of the 8 step lengths (glass::reduce over the knots, as
// include/common/merit.cuh:86-93), then the decision and update of include/pcg/sqp.cuh:352-413.
template <typename T>
__global__ void select_step_batched(uint32_t knot_points, uint32_t nz, T* d_point, const T* d_dz, T* d_xu, size_t xu_stride,
                                    BatchedSqpState<T> st, T rho_factor, T rho_min, T rho_max, T rho_reset) {
    const size_t b = blockIdx.x;
    __shared__ T s_merit[kLsAlphas];
    __shared__ T s_alpha;
    for (uint32_t p = 0; p < kLsAlphas; p++) {
        T* pm = d_point + (b * (kLsAlphas + 1) + p) * knot_points;
        glass::reduce<T>(knot_points, pm);
        if (threadIdx.x == 0) s_merit[p] = pm[0];
    }
    __syncthreads();
    if (threadIdx.x == 0 && (st.flags[b] & 1u)) {      // left the loop (rho ran away): frozen, like the reference's `break`
        s_alpha = T(0);
        st.step[b] = T(0);
    } else if (threadIdx.x == 0) {
        const T merit_now = st.merit[b];
        uint32_t best = 0;
        T best_merit = merit_now;
        for (uint32_t p = 0; p < kLsAlphas; p++)
            if (s_merit[p] < best_merit) { best_merit = s_merit[p]; best = p; }
        T rho = st.rho[b], drho = st.drho[b];
        if (best_merit == merit_now) {        // no step length improves: raise the regularisation (:372-384)
            drho = max(drho * rho_factor, rho_factor);
            rho = max(rho * drho, rho_min);
            if (rho > rho_max) { rho = rho_reset; st.flags[b] |= 1u; }
            s_alpha = T(0);
        } else {                              // accept (:388-413)
            drho = min(drho / rho_factor, T(1) / rho_factor);
            rho = max(rho * drho, rho_min);
            s_alpha = T(-1.0) / T(1u << best);
            st.merit[b] = best_merit;
        }
        st.rho[b] = rho; st.drho[b] = drho; st.step[b] = s_alpha;
    }
    __syncthreads();
    const T alpha = s_alpha;
    if (alpha != T(0)) {
        T* xu = d_xu + b * xu_stride;
        const T* dz = d_dz + b * xu_stride;
        for (uint32_t i = threadIdx.x; i < nz; i += blockDim.x) xu[i] += alpha * dz[i];
    }
}

// Device workspace of `batch` instances (caller-owned lifetime; nothing is allocated per iteration).
template <typename T>
struct BatchedSqp {
    uint32_t s = 0, c = 0, N = 0, B = 0, nz = 0;
    T *G = nullptr, *C = nullptr, *g = nullptr, *cv = nullptr, *dz = nullptr, *point = nullptr;
    uint32_t* iters = nullptr;
    uint8_t* exits = nullptr;
    BatchedSqpState<T> st{};

    cudaError_t create(uint32_t s_, uint32_t c_, uint32_t N_, uint32_t B_, T rho0) {
        static_assert(std::is_same<T, float>::value, "BatchedSqp: the per-instance-rho solve exists in fp32 only "
                                                     "(mpcgpu_b200_kkt_solve_dz_rhovec_f32); build with USE_DOUBLES 0");
        if (B_ == 0 || B_ > 65535 || N_ < 2) return cudaErrorInvalidValue;     // the batch index is gridDim.y / gridDim.z
        s = s_; c = c_; N = N_; B = B_; nz = (s + c) * N - c;
        const size_t zG = (size_t)(s * s + c * c) * N - c * c, zC = (size_t)(s * s + s * c) * (N - 1);
        cudaError_t e;
        if ((e = cudaMalloc(&G, B * zG * sizeof(T))) != cudaSuccess) return e;
        if ((e = cudaMalloc(&C, B * zC * sizeof(T))) != cudaSuccess) return e;
        if ((e = cudaMalloc(&g, (size_t)B * nz * sizeof(T))) != cudaSuccess) return e;
        if ((e = cudaMalloc(&cv, (size_t)B * s * N * sizeof(T))) != cudaSuccess) return e;
        if ((e = cudaMalloc(&dz, (size_t)B * nz * sizeof(T))) != cudaSuccess) return e;
        if ((e = cudaMalloc(&point, (size_t)B * (kLsAlphas + 1) * N * sizeof(T))) != cudaSuccess) return e;
        if ((e = cudaMalloc(&iters, B * sizeof(uint32_t))) != cudaSuccess) return e;
        if ((e = cudaMalloc(&exits, B)) != cudaSuccess) return e;
        if ((e = cudaMalloc(&st.rho, B * sizeof(T))) != cudaSuccess) return e;
        if ((e = cudaMalloc(&st.drho, B * sizeof(T))) != cudaSuccess) return e;
        if ((e = cudaMalloc(&st.merit, B * sizeof(T))) != cudaSuccess) return e;
        if ((e = cudaMalloc(&st.step, B * sizeof(T))) != cudaSuccess) return e;
        if ((e = cudaMalloc(&st.flags, B * sizeof(uint32_t))) != cudaSuccess) return e;
        std::vector<T> r(B, rho0), d1(B, T(1));
        if ((e = cudaMemcpy(st.rho, r.data(), B * sizeof(T), cudaMemcpyHostToDevice)) != cudaSuccess) return e;
        if ((e = cudaMemcpy(st.drho, d1.data(), B * sizeof(T), cudaMemcpyHostToDevice)) != cudaSuccess) return e;
        if ((e = cudaMemset(st.flags, 0, B * sizeof(uint32_t))) != cudaSuccess) return e;
        return cudaMemset(cv, 0, (size_t)B * s * N * sizeof(T));
    }
    void destroy() {
        for (void* p : {(void*)G, (void*)C, (void*)g, (void*)cv, (void*)dz, (void*)point, (void*)iters, (void*)exits,
                        (void*)st.rho, (void*)st.drho, (void*)st.merit, (void*)st.step, (void*)st.flags})
            cudaFree(p);
        G = nullptr;
    }
    size_t merit_smem() const { return get_merit_smem_size<T>(s, N > c ? N : c); }   // the reference's size (include/pcg/sqp.cuh:343)

    // merit of the starting iterates (include/pcg/sqp.cuh:186-193), fixed summation order
    cudaError_t init_merit(const T* d_xs, const T* d_xu, const T* d_ee, size_t ee_stride, void* d_dyn, T dt, T mu,
                           cudaStream_t stream) {
        ls_merit_batched<T><<<dim3(N, 1, B), batched_threads("B200SLOT_MERIT_THREADS", 32), merit_smem(), stream>>>(s, c, N, d_xs, d_xu, nz, d_ee, ee_stride,
                                                                                  mu, dt, d_dyn, dz, kLsAlphas, point);
        init_merit_batched<T><<<B, MERIT_THREADS, 0, stream>>>(N, point, st.merit);
        return cudaGetLastError();
    }

    // one SQP iteration of every instance: four launches on `stream`, no host synchronisation
    cudaError_t iterate(T* d_xu, const T* d_xs, const T* d_ee, size_t ee_stride, T* d_lambda, void* d_dyn, T dt, T mu,
                        uint32_t pcg_max_iter, T pcg_exit_tol, T rho_reset, cudaStream_t stream) {
        cudaError_t e = launch_kkt_batched<T>(s, c, N, B, G, C, g, cv, d_dyn, dt, d_ee, ee_stride, d_xs, d_xu, nz, stream);
        if (e != cudaSuccess) return e;
        e = (cudaError_t)mpcgpu_b200_kkt_solve_dz_rhovec_f32(s, c, N, B, G, C, g, cv, st.rho, d_lambda, dz, pcg_max_iter,
                                                            pcg_exit_tol, iters, exits, nullptr, (void*)stream);
        if (e != cudaSuccess) return e;
        ls_merit_batched<T><<<dim3(N, kLsAlphas, B), batched_threads("B200SLOT_MERIT_THREADS", 32), merit_smem(), stream>>>(
            s, c, N, d_xs, d_xu, nz, d_ee, ee_stride, mu, dt, d_dyn, dz, 0, point);
        select_step_batched<T><<<B, MERIT_THREADS, 0, stream>>>(N, nz, point, dz, d_xu, nz, st, T(RHO_FACTOR), T(RHO_MIN),
                                                              T(RHO_MAX), rho_reset);
        return cudaGetLastError();
    }
};

}  // namespace b200slot
