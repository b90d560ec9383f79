// b200/sqp.cuh -- third linear-solver slot for MPCGPU ("LINSYS_SOLVE == 2"), next to
// include/pcg/sqp.cuh (sqpSolvePcg) and include/qdldl/sqp.cuh (sqpSolveQdldl).
//
//   sqpSolveB200<T>(state_size, control_size, knot_points, timestep, d_eePos_traj, d_lambda,
//                   d_xu, d_dynMem_const, pcg_config<T>&, T& rho, T rho_reset)
//
// Same arguments, same in/out buffers (d_lambda warm start, d_xu updated in place, rho adapted)
// and the same 10-tuple as include/pcg/sqp.cuh:26,463-474, so include/mpcsim.cuh:249-277,303-307
// can call it unchanged.  What differs from the PCG slot:
//   * form_schur_system + cooperative pcg launch + two blocking read-backs + compute_dz
//     (include/pcg/sqp.cuh:243-309) are ONE launch of mpcgpu_b200_kkt_solve_dz_f32; S and Pinv are
//     never materialised in HBM, the iteration count comes home asynchronously;
//   * every buffer, stream and pinned mailbox lives in a workspace built on the first call and
//     kept (the reference slot does ~20 cudaMalloc/cudaFree, 8 stream creates and a cuBLAS
//     handle per SQP solve, inside its own 2 ms budget);
//   * the accepted step xu += alpha dz is a 1-block kernel instead of a cuBLAS call;
//   * one SQP iteration (step, KKT blocks, solve + dz, the 8 line-search kernels on 8 streams, the
//     read-backs) is captured ONCE into a CUDA graph and replayed: one graph launch and one stream
//     synchronise per iteration; rho and the step length enter through a pinned 8-byte control word
//     (MPCGPU_B200_SLOT_GRAPH=0, FINE_GRAINED_TIMING or TIME_LINSYS select the plain stream path).
// The dynamics-facing kernels are the reference's own, included from where they are:
// generate_kkt_submatrices (common/kkt.cuh), compute_merit and ls_gato_compute_merit
// (common/merit.cuh).  linsys_t = float or double (settings.cuh USE_DOUBLES; double: knot_points <= 256).
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <tuple>
#include <vector>
#include <type_traits>
#include <time.h>
#include <cuda_runtime.h>
#include "settings.cuh"
#include "gpuassert.cuh"
#include "types.cuh"
#ifndef MPCGPU_B200_SLOT_COMMON_INCLUDED   // common/kkt.cuh and common/dz.cuh carry no include guard
#include "common/kkt.cuh"
#include "common/dz.cuh"
#endif
#include "merit.cuh"
#include "mpcgpu_b200.h"

namespace b200slot {

constexpr uint32_t kAlphas = 8;       // step lengths 1, 1/2, ... 1/128 (include/pcg/sqp.cuh:63,410)
constexpr float    kMu     = 10.0f;   // merit penalty (include/pcg/sqp.cuh:62,192)

inline double us_between(const timespec& a, const timespec& b) {
    return 1e6 * double(b.tv_sec - a.tv_sec) + 1e-3 * double(b.tv_nsec - a.tv_nsec);
}

// linsys_t = float / double (settings.cuh USE_DOUBLES) -> the _f32 / _f64 entry points of the C ABI
inline int solve_dz(uint32_t s, uint32_t c, uint32_t N, uint32_t B, float* G, const float* C, const float* g, const float* cv,
                    float rho, float* lam, float* dz, uint32_t mi, float tol, uint32_t* it, uint8_t* ex, void* ws, void* st) {
    return mpcgpu_b200_kkt_solve_dz_f32(s, c, N, B, G, C, g, cv, rho, lam, dz, mi, tol, it, ex, ws, st);
}
inline int solve_dz(uint32_t s, uint32_t c, uint32_t N, uint32_t B, double* G, const double* C, const double* g, const double* cv,
                    double rho, double* lam, double* dz, uint32_t mi, double tol, uint32_t* it, uint8_t* ex, void* ws, void* st) {
    return mpcgpu_b200_kkt_solve_dz_f64(s, c, N, B, G, C, g, cv, rho, lam, dz, mi, tol, it, ex, ws, st);
}
inline int solve_dz_dev(uint32_t s, uint32_t c, uint32_t N, uint32_t B, float* G, const float* C, const float* g, const float* cv,
                        const float* rho, float* lam, float* dz, uint32_t mi, float tol, uint32_t* it, uint8_t* ex, void* ws, void* st) {
    return mpcgpu_b200_kkt_solve_dz_dev_f32(s, c, N, B, G, C, g, cv, rho, lam, dz, mi, tol, it, ex, ws, st);
}
inline int solve_dz_dev(uint32_t s, uint32_t c, uint32_t N, uint32_t B, double* G, const double* C, const double* g, const double* cv,
                        const double* rho, double* lam, double* dz, uint32_t mi, double tol, uint32_t* it, uint8_t* ex, void* ws, void* st) {
    return mpcgpu_b200_kkt_solve_dz_dev_f64(s, c, N, B, G, C, g, cv, rho, lam, dz, mi, tol, it, ex, ws, st);
}

template <typename T>
__global__ void take_step(uint32_t n, T alpha, const T* __restrict__ dz, T* __restrict__ xu) {
    for (uint32_t i = threadIdx.x + blockIdx.x * blockDim.x; i < n; i += blockDim.x * gridDim.x)
        xu[i] += alpha * dz[i];
}

// graph replay: the step length lives in device memory; 0 = the last line search found no step
template <typename T>
__global__ void take_step_dev(uint32_t n, const T* __restrict__ alpha_dev, const T* __restrict__ dz, T* __restrict__ xu) {
    const T alpha = *alpha_dev;
    if (alpha == T(0)) return;
    for (uint32_t i = threadIdx.x + blockIdx.x * blockDim.x; i < n; i += blockDim.x * gridDim.x)
        xu[i] += alpha * dz[i];
}

// What one (state_size, control_size, knot_points) needs on the device, carved out of a single
// allocation; pinned host words for everything the host reads back.
template <typename T>
struct Workspace {
    uint32_t s = 0, c = 0, N = 0;
    char* d_pool = nullptr;
    T *G = nullptr, *C = nullptr, *g = nullptr, *cv = nullptr, *dz = nullptr, *xs = nullptr;
    T *merit0 = nullptr, *merits = nullptr, *merit_tmp = nullptr;
    uint32_t* iters = nullptr;   // [SQP_MAX_ITER]
    uint8_t*  flags = nullptr;   // [SQP_MAX_ITER]
    struct Host { T merit0; T merits[kAlphas]; uint32_t iters[SQP_MAX_ITER]; uint8_t flags[SQP_MAX_ITER]; T ctl[2]; };
    T* ctl = nullptr;            // device copy of Host::ctl = {rho, step length}
    // the captured SQP iteration and the arguments it was captured for
    cudaStream_t main = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join[kAlphas] = {};
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    bool graph_failed = false;
    const void* key_ptr[4] = {nullptr, nullptr, nullptr, nullptr};
    float key_dt = 0.f;
    T key_tol = T(0);            // as T: a double tolerance (REMOVE_JITTERS sets 1e-11) must compare equal to itself
    uint32_t key_iter = 0;
    int device = -1;             // the device the pool lives on
    Host* h = nullptr;
    cudaStream_t ls_stream[kAlphas] = {};
    uint32_t nz = 0;             // length of xu / dz

    static size_t up(size_t x) { return (x + 255) & ~size_t(255); }

    void build(uint32_t s_, uint32_t c_, uint32_t N_) {
        release();
        s = s_; c = c_; N = N_;
        gpuErrchk(cudaGetDevice(&device));
        nz = (s + c) * N - c;
        const size_t nG = size_t(s * s + c * c) * N - c * c, nC = size_t(s * s + s * c) * (N - 1),
                     ng = nz, nc = size_t(s) * N;
        const size_t sz[] = {nG * sizeof(T), nC * sizeof(T), ng * sizeof(T), nc * sizeof(T), size_t(nz) * sizeof(T),
                             s * sizeof(T), sizeof(T), kAlphas * sizeof(T), size_t(kAlphas) * N * sizeof(T),
                             SQP_MAX_ITER * sizeof(uint32_t), SQP_MAX_ITER * sizeof(uint8_t), 2 * sizeof(T)};
        size_t off[13] = {0};
        for (int i = 0; i < 12; i++) off[i + 1] = off[i] + up(sz[i]);
        gpuErrchk(cudaMalloc(&d_pool, off[12]));
        gpuErrchk(cudaMemset(d_pool, 0, off[12]));
        ctl = (T*)(d_pool + off[11]);
        G = (T*)(d_pool + off[0]);  C = (T*)(d_pool + off[1]);  g = (T*)(d_pool + off[2]);
        cv = (T*)(d_pool + off[3]); dz = (T*)(d_pool + off[4]); xs = (T*)(d_pool + off[5]);
        merit0 = (T*)(d_pool + off[6]); merits = (T*)(d_pool + off[7]); merit_tmp = (T*)(d_pool + off[8]);
        iters = (uint32_t*)(d_pool + off[9]); flags = (uint8_t*)(d_pool + off[10]);
        gpuErrchk(cudaMallocHost(&h, sizeof(Host)));
        for (uint32_t i = 0; i < kAlphas; i++) gpuErrchk(cudaStreamCreate(&ls_stream[i]));
        // blocking streams: ordered against the caller's legacy-stream work like the reference's own
        gpuErrchk(cudaStreamCreate(&main));
        gpuErrchk(cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming));
        for (uint32_t i = 0; i < kAlphas; i++) gpuErrchk(cudaEventCreateWithFlags(&ev_join[i], cudaEventDisableTiming));
    }
    void drop_graph() {
        if (exec) cudaGraphExecDestroy(exec);
        if (graph) cudaGraphDestroy(graph);
        exec = nullptr; graph = nullptr;
    }
    void release() {
        if (!d_pool) return;
        drop_graph();
        for (uint32_t i = 0; i < kAlphas; i++) { cudaStreamDestroy(ls_stream[i]); cudaEventDestroy(ev_join[i]); }
        cudaEventDestroy(ev_fork); cudaStreamDestroy(main);
        graph_failed = false;
        cudaFreeHost(h); cudaFree(d_pool);
        d_pool = nullptr; h = nullptr;
    }
    ~Workspace() { /* process-lifetime object: the driver reclaims the memory at exit */ }
};

// One workspace per process, rebuilt when the problem size or the current device changes.  Like the reference's
// slots (one host thread, legacy stream: SURVEY 8b "Threading / streams") it is NOT thread-safe: call
// sqpSolveB200 from one host thread at a time.
template <typename T>
inline Workspace<T>& workspace(uint32_t s, uint32_t c, uint32_t N) {
    static Workspace<T> w;
    int dev = -1;
    gpuErrchk(cudaGetDevice(&dev));
    if (w.s != s || w.c != c || w.N != N || !w.d_pool || w.device != dev) w.build(s, c, N);
    return w;
}

}  // namespace b200slot

template <typename T>
auto sqpSolveB200(const uint32_t state_size, const uint32_t control_size, const uint32_t knot_points, float timestep,
                  T* d_eePos_traj, T* d_lambda, T* d_xu, void* d_dynMem_const, pcg_config<T>& config, T& rho,
                  T rho_reset) {
    static_assert(std::is_same<T, float>::value || std::is_same<T, double>::value, "mpcgpu_b200 slot: linsys_t is float or double");
    using namespace b200slot;

    std::vector<int> pcg_iter_vec;
    std::vector<bool> pcg_exit_vec;
    std::vector<double> linsys_time_vec, kkt_time_vec, schur_time_vec, dz_time_vec, line_search_time_vec;
    bool sqp_time_exit = true;   // false only when rho runs away (include/pcg/sqp.cuh:395-399)
    uint32_t sqp_iter = 0, solves = 0;

    timespec t_begin, t_now, t_a, t_b;
    gpuErrchk(cudaDeviceSynchronize());
    clock_gettime(CLOCK_MONOTONIC, &t_begin);
    auto out_of_time = [&]() -> bool {
#if CONST_UPDATE_FREQ
        clock_gettime(CLOCK_MONOTONIC, &t_now);
        return us_between(t_begin, t_now) > SQP_MAX_TIME_US;
#else
        return false;
#endif
    };
    // instrumentation brackets (blocking, like the reference's): compiled out unless asked for
    auto lap_begin = [&](bool on) { if (on) { gpuErrchk(cudaDeviceSynchronize()); clock_gettime(CLOCK_MONOTONIC, &t_a); } };
    auto lap_end = [&](bool on, std::vector<double>& v) {
        if (on) { gpuErrchk(cudaDeviceSynchronize()); clock_gettime(CLOCK_MONOTONIC, &t_b); v.push_back(us_between(t_a, t_b)); }
    };
    constexpr bool fine = FINE_GRAINED_TIMING != 0, time_linsys = TIME_LINSYS != 0;

    Workspace<T>& w = workspace<T>(state_size, control_size, knot_points);
    gpuErrchk(cudaMemcpyAsync(w.xs, d_xu, state_size * sizeof(T), cudaMemcpyDeviceToDevice));
    gpuErrchk(cudaMemsetAsync(w.merit0, 0, sizeof(T)));

    const size_t merit_smem = get_merit_smem_size<T>(state_size, control_size);
    const size_t ls_smem = get_merit_smem_size<T>(state_size, knot_points);   // as include/pcg/sqp.cuh:356
    const size_t kkt_smem = 2 * get_kkt_smem_size<T>(state_size, control_size);
    void* ls_kernel = (void*)ls_gato_compute_merit<T>;
    uint32_t ks = state_size, kc = control_size, kn = knot_points;
    T mu = kMu, dt = timestep;

    compute_merit<T><<<knot_points, MERIT_THREADS, merit_smem>>>(state_size, control_size, knot_points, d_xu,
                                                                  d_eePos_traj, mu, dt, d_dynMem_const, w.merit0);
    gpuErrchk(cudaPeekAtLastError());
    gpuErrchk(cudaMemcpyAsync(&w.h->merit0, w.merit0, sizeof(T), cudaMemcpyDeviceToHost));
    bool merit0_home = false;
    T merit_now = 0, drho = 1;
    const T rho_factor = RHO_FACTOR, rho_max = RHO_MAX, rho_min = RHO_MIN;

    // MPCGPU_B200_SLOT_MAX_SQP=n caps the SQP iterations of a call below SQP_MAX_ITER (closed-loop experiments:
    // the b200 slot with the iteration count the reference's PCG slot reaches in the same budget)
    static const uint32_t sqp_cap = [] {
        const char* e = getenv("MPCGPU_B200_SLOT_MAX_SQP");
        const long v = e ? atol(e) : 0;
        return (v > 0 && v < (long)SQP_MAX_ITER) ? (uint32_t)v : (uint32_t)SQP_MAX_ITER;
    }();
    // ---- graph path: one captured SQP iteration, replayed ---------------------------------------
    static const bool want_graph = [] { const char* e = getenv("MPCGPU_B200_SLOT_GRAPH"); return !(e && e[0] == '0'); }();
    bool use_graph = want_graph && !fine && !time_linsys && !w.graph_failed;
    if (use_graph) {
        const void* key[4] = {d_eePos_traj, d_lambda, d_xu, d_dynMem_const};
        const bool same = w.exec && key[0] == w.key_ptr[0] && key[1] == w.key_ptr[1] && key[2] == w.key_ptr[2] &&
                          key[3] == w.key_ptr[3] && w.key_dt == timestep && w.key_tol == config.pcg_exit_tol &&
                          w.key_iter == config.pcg_max_iter;
        if (!same) {
            w.drop_graph();
            bool ok = cudaStreamBeginCapture(w.main, cudaStreamCaptureModeThreadLocal) == cudaSuccess;
            if (ok) {
                ok &= cudaMemcpyAsync(w.ctl, w.h->ctl, 2 * sizeof(T), cudaMemcpyHostToDevice, w.main) == cudaSuccess;
                take_step_dev<T><<<1, 256, 0, w.main>>>(w.nz, w.ctl + 1, w.dz, d_xu);
                generate_kkt_submatrices<T><<<knot_points, KKT_THREADS, kkt_smem, w.main>>>(
                    state_size, control_size, knot_points, w.G, w.C, w.g, w.cv, d_dynMem_const, dt, d_eePos_traj, w.xs, d_xu);
                ok &= solve_dz_dev(state_size, control_size, knot_points, 1, w.G, w.C, w.g, w.cv, w.ctl,
                                                       d_lambda, w.dz, config.pcg_max_iter, config.pcg_exit_tol, w.iters,
                                                       w.flags, nullptr, (void*)w.main) == 0;
                ok &= cudaEventRecord(w.ev_fork, w.main) == cudaSuccess;
                for (uint32_t p = 0; p < kAlphas; p++) {
                    void* args[] = {&ks, &kc, &kn, &w.xs, &d_xu, &d_eePos_traj, &mu, &dt, &d_dynMem_const, &w.dz, &p,
                                    &w.merits, &w.merit_tmp};
                    ok &= cudaStreamWaitEvent(w.ls_stream[p], w.ev_fork, 0) == cudaSuccess;
                    ok &= cudaLaunchCooperativeKernel(ls_kernel, knot_points, MERIT_THREADS, args, ls_smem, w.ls_stream[p]) == cudaSuccess;
                    ok &= cudaEventRecord(w.ev_join[p], w.ls_stream[p]) == cudaSuccess;
                    ok &= cudaStreamWaitEvent(w.main, w.ev_join[p], 0) == cudaSuccess;
                }
                ok &= cudaMemcpyAsync(w.h->merits, w.merits, kAlphas * sizeof(T), cudaMemcpyDeviceToHost, w.main) == cudaSuccess;
                ok &= cudaMemcpyAsync(&w.h->iters[0], w.iters, sizeof(uint32_t), cudaMemcpyDeviceToHost, w.main) == cudaSuccess;
                ok &= cudaMemcpyAsync(&w.h->flags[0], w.flags, sizeof(uint8_t), cudaMemcpyDeviceToHost, w.main) == cudaSuccess;
                ok &= cudaStreamEndCapture(w.main, &w.graph) == cudaSuccess && w.graph != nullptr;
                ok = ok && cudaGraphInstantiate(&w.exec, w.graph, 0) == cudaSuccess;
            }
            if (!ok) {   // capture not possible here (e.g. cooperative launches): plain streams from now on
                cudaGetLastError();
                w.drop_graph();
                w.graph_failed = true;
                use_graph = false;
            } else {
                for (int i = 0; i < 4; i++) w.key_ptr[i] = key[i];
                w.key_dt = timestep; w.key_tol = config.pcg_exit_tol; w.key_iter = config.pcg_max_iter;
            }
        }
    }
    if (use_graph) {
        T pending = 0;                       // accepted step, applied by the next replay (or after the loop)
        double iter_us = 0.0;                // duration of the last replay: do not start one that cannot finish
        for (uint32_t it = 0; it < sqp_cap; it++) {
#if CONST_UPDATE_FREQ
            clock_gettime(CLOCK_MONOTONIC, &t_now);
            if (us_between(t_begin, t_now) + iter_us > SQP_MAX_TIME_US) break;
#endif
            timespec r0, r1;
            clock_gettime(CLOCK_MONOTONIC, &r0);
            w.h->ctl[0] = rho; w.h->ctl[1] = pending; pending = 0;
            gpuErrchk(cudaGraphLaunch(w.exec, w.main));
            gpuErrchk(cudaStreamSynchronize(w.main));
            clock_gettime(CLOCK_MONOTONIC, &r1);
            iter_us = us_between(r0, r1);
            pcg_iter_vec.push_back(int(w.h->iters[0]));
            pcg_exit_vec.push_back(w.h->flags[0] != 0);
            if (!merit0_home) { merit_now = w.h->merit0; merit0_home = true; }
            uint32_t best = 0;
            T best_merit = merit_now;
            for (uint32_t i = 0; i < kAlphas; i++)
                if (w.h->merits[i] < best_merit) { best_merit = w.h->merits[i]; best = i; }
            if (best_merit == merit_now) {
                drho = max(drho * rho_factor, rho_factor);
                rho = max(rho * drho, rho_min);
                sqp_iter++;
                if (rho > rho_max) { sqp_time_exit = false; rho = rho_reset; break; }
                continue;
            }
            drho = min(drho / rho_factor, 1 / rho_factor);
            rho = max(rho * drho, rho_min);
            pending = T(-1.0) / T(1u << best);
            sqp_iter++;
            merit_now = best_merit;
        }
        if (pending != T(0)) {
            take_step<T><<<1, 256, 0, w.main>>>(w.nz, pending, w.dz, d_xu);
            gpuErrchk(cudaPeekAtLastError());
        }
    }

    for (uint32_t it = 0; !use_graph && it < sqp_cap; it++) {
        // 1. linearise: G, C, g, c
        lap_begin(fine);
        generate_kkt_submatrices<T><<<knot_points, KKT_THREADS, kkt_smem>>>(state_size, control_size, knot_points, w.G,
                                                                            w.C, w.g, w.cv, d_dynMem_const, dt,
                                                                            d_eePos_traj, w.xs, d_xu);
        gpuErrchk(cudaPeekAtLastError());
        lap_end(fine, kkt_time_vec);
        if (out_of_time()) break;

        // 2.+3. Schur complement + preconditioner + PCG + dz recovery: one launch; lambda warm-started
        //    in place, G becomes G^-1 in place, dz = compute_dz(G^-1, C, g, lambda) as the kernel's epilogue
        lap_begin(fine || time_linsys);
        gpuErrchk((cudaError_t)solve_dz(state_size, control_size, knot_points, 1, w.G, w.C, w.g,
                                                            w.cv, rho, d_lambda, w.dz, config.pcg_max_iter,
                                                            config.pcg_exit_tol, w.iters + solves, w.flags + solves,
                                                            nullptr, nullptr));
        gpuErrchk(cudaMemcpyAsync(&w.h->iters[solves], w.iters + solves, sizeof(uint32_t), cudaMemcpyDeviceToHost));
        gpuErrchk(cudaMemcpyAsync(&w.h->flags[solves], w.flags + solves, sizeof(uint8_t), cudaMemcpyDeviceToHost));
        solves++;
        lap_end(time_linsys, linsys_time_vec);
        if (fine) { schur_time_vec.push_back(0.0); dz_time_vec.push_back(0.0); }   // both inside the solve launch
        if (out_of_time()) break;

        // 4. parallel line search over 8 step lengths, one stream each
        lap_begin(fine);
        for (uint32_t p = 0; p < kAlphas; p++) {
            void* args[] = {&ks, &kc, &kn, &w.xs, &d_xu, &d_eePos_traj, &mu, &dt, &d_dynMem_const, &w.dz, &p,
                            &w.merits, &w.merit_tmp};
            gpuErrchk(cudaLaunchCooperativeKernel(ls_kernel, knot_points, MERIT_THREADS, args, ls_smem, w.ls_stream[p]));
        }
        gpuErrchk(cudaPeekAtLastError());
        lap_end(fine, line_search_time_vec);
        if (out_of_time()) break;
        // blocking copy on the legacy stream: waits for the 8 streams, the solve and its read-backs
        gpuErrchk(cudaMemcpy(w.h->merits, w.merits, kAlphas * sizeof(T), cudaMemcpyDeviceToHost));
        if (!merit0_home) { merit_now = w.h->merit0; merit0_home = true; }
        if (out_of_time()) break;

        uint32_t best = 0;
        T best_merit = merit_now;
        for (uint32_t i = 0; i < kAlphas; i++)
            if (w.h->merits[i] < best_merit) { best_merit = w.h->merits[i]; best = i; }

        if (best_merit == merit_now) {   // no step length improves: raise the regularisation
            drho = max(drho * rho_factor, rho_factor);
            rho = max(rho * drho, rho_min);
            sqp_iter++;
            if (rho > rho_max) { sqp_time_exit = false; rho = rho_reset; break; }
            continue;
        }
        drho = min(drho / rho_factor, 1 / rho_factor);
        rho = max(rho * drho, rho_min);
        take_step<T><<<1, 256>>>(w.nz, T(-1.0) / T(1u << best), w.dz, d_xu);
        gpuErrchk(cudaPeekAtLastError());
        sqp_iter++;
        if (out_of_time()) break;
        merit_now = best_merit;
    }

    gpuErrchk(cudaPeekAtLastError());
    gpuErrchk(cudaDeviceSynchronize());   // also lands every asynchronous read-back
    timespec t_end;
    clock_gettime(CLOCK_MONOTONIC, &t_end);
    for (uint32_t i = 0; i < solves; i++) {          // stream path: the asynchronous read-backs have landed
        pcg_iter_vec.push_back(int(w.h->iters[i]));
        pcg_exit_vec.push_back(w.h->flags[i] != 0);
    }
    const double sqp_solve_time = us_between(t_begin, t_end);
    return std::make_tuple(pcg_iter_vec, linsys_time_vec, sqp_solve_time, sqp_iter, sqp_time_exit, pcg_exit_vec,
                           kkt_time_vec, schur_time_vec, dz_time_vec, line_search_time_vec);
}
