// e2e_producer.cu -- BASELINE.json configs[4] END TO END with the KKT producer on the device (VERDICT r1 "Next" #5):
// per step the host ships only what an MPC instance really owns -- its iterate xu ((s+c)N - c floats), its measured
// start state xs (s) and its end-effector goal window (6N): ~7 KB per 64-knot instance instead of the 149 KB of
// KKT blocks -- and gets back the Newton step dz:
//     pinned host buffers --H2D--> b200slot::generate_kkt_submatrices_batched (the reference's own GRiD dynamics and
//     cost device functions, one launch over (knot, instance)) --> mpcgpu_b200_kkt_solve_dz_f32 (Schur + PCG +
//     compute_dz, one launch) --D2H--> dz, iteration counts
// Timed region: `steps` such steps on one stream, host clock around them with a synchronise on both sides.
// One JSON line.  Needs the reference's headers to build (oracle/build.py where /root/reference exists); the
// binary lives in mpcgpu_b200/_plugins/ and reads examples/trajfiles/ relative to its working directory.
//   e2e_producer [N=64] [B=4096] [tol=1e-4] [steps=10] [warmup=3] [device=0]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <string>
#include <vector>
#include <cstdint>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <sstream>
#include <tuple>
#include <time.h>
#include <cooperative_groups.h>

#include "settings.cuh"
#define MPCGPU_B200_SHIM_KEEP_REFERENCE_SCHUR 1
#include "gpu_pcg.cuh"
#include "common/kkt.cuh"
#include "utils/experiment.cuh"
#define MPCGPU_B200_SLOT_COMMON_INCLUDED 1
#include "b200/kkt_batched.cuh"
#include "mpcgpu_b200.h"

#define CK(x) gpuErrchk(x)
template <typename T> static T* dalloc(size_t n) { T* d; CK(cudaMalloc(&d, std::max<size_t>(n, 1) * sizeof(T))); return d; }
template <typename T> static T* halloc(size_t n) { T* h; CK(cudaHostAlloc((void**)&h, std::max<size_t>(n, 1) * sizeof(T), cudaHostAllocDefault)); return h; }
static double now_s() { timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return double(t.tv_sec) + 1e-9 * double(t.tv_nsec); }

int main(int argc, char** argv) {
    const uint32_t s = 14, m = 7;
    const uint32_t N = argc > 1 ? (uint32_t)atoi(argv[1]) : 64;
    const uint32_t B = argc > 2 ? (uint32_t)atoi(argv[2]) : 4096;
    const float tol = argc > 3 ? (float)atof(argv[3]) : 1e-4f;
    const int steps = argc > 4 ? atoi(argv[4]) : 10, warmup = argc > 5 ? atoi(argv[5]) : 3;
    const int device = argc > 6 ? atoi(argv[6]) : 0;
    CK(cudaSetDevice(device));
    const float rho = 1e-3f, timestep = 0.015625f;          // include/mpcsim.cuh:256, examples/track_iiwa_pcg.cu:19
    auto ee = readCSVToVecVec<float>("examples/trajfiles/0_0_eepos.traj");
    auto xu = readCSVToVecVec<float>("examples/trajfiles/0_0_traj.csv");
    if (ee.size() < N + 8 || xu.size() < N + 8) { fprintf(stderr, "trajectory too short\n"); return 3; }
    const size_t zG = (size_t)(s * s + m * m) * N - m * m, zC = (size_t)(s * s + s * m) * (N - 1),
                 zg = (size_t)(s + m) * N - m, zc = (size_t)s * N, zv = (size_t)s * N, zee = 6 * (size_t)N;
    // instance b: a window of the nominal trajectory, iterate perturbed by a deterministic ripple, start state off x_0
    float *h_xu = halloc<float>(B * zg), *h_ee = halloc<float>(B * zee), *h_xs = halloc<float>(B * (size_t)s);
    float* h_dz = halloc<float>(B * zg);
    uint32_t* h_it = halloc<uint32_t>(B);
    uint8_t* h_ex = halloc<uint8_t>(B);
    for (uint32_t b = 0; b < B; b++) {
        const uint32_t row0 = b % 8;
        const float amp = 0.002f + 0.02f * (float)b / (float)B;
        size_t o = 0;
        for (uint32_t k = 0; k < N; k++) {
            for (int i = 0; i < 6; i++) h_ee[b * zee + 6 * k + i] = ee[row0 + k][i];
            const size_t take = (k + 1 < N) ? s + m : s;
            for (size_t i = 0; i < take; i++, o++)
                h_xu[b * zg + o] = xu[row0 + k][i] + amp * std::sin(0.37f * (float)o + 1.3f + 0.11f * (float)b);
        }
        for (uint32_t i = 0; i < s; i++) h_xs[b * s + i] = h_xu[b * zg + i] + 0.5f * amp * std::cos(0.9f * (float)i + (float)b);
    }
    float *d_xu = dalloc<float>(B * zg), *d_ee = dalloc<float>(B * zee), *d_xs = dalloc<float>(B * (size_t)s);
    float *d_G = dalloc<float>(B * zG), *d_C = dalloc<float>(B * zC), *d_g = dalloc<float>(B * zg), *d_c = dalloc<float>(B * zc);
    float *d_lam = dalloc<float>(B * zv), *d_dz = dalloc<float>(B * zg);
    uint32_t* d_it = dalloc<uint32_t>(B);
    uint8_t* d_ex = dalloc<uint8_t>(B);
    void* d_dyn = gato_plant::initializeDynamicsConstMem<float>();
    cudaStream_t st;
    CK(cudaStreamCreate(&st));
    cudaEvent_t e0, e1, e2, e3;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1)); CK(cudaEventCreate(&e2)); CK(cudaEventCreate(&e3));
    double ms_h2d = 0, ms_kkt = 0, ms_solve = 0, t_begin = 0;
    const uint64_t l0 = mpcgpu_b200_launch_count();
    for (int r = -warmup; r < steps; r++) {
        if (r == 0) { CK(cudaDeviceSynchronize()); t_begin = now_s(); }
        CK(cudaEventRecord(e0, st));
        CK(cudaMemcpyAsync(d_xu, h_xu, B * zg * 4, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(d_ee, h_ee, B * zee * 4, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(d_xs, h_xs, B * (size_t)s * 4, cudaMemcpyHostToDevice, st));
        CK(cudaMemsetAsync(d_lam, 0, B * zv * 4, st));                                  // cold start, like bench.py's e2e
        CK(cudaEventRecord(e1, st));
        CK(b200slot::launch_kkt_batched<float>(s, m, N, B, d_G, d_C, d_g, d_c, d_dyn, timestep, d_ee, zee, d_xs, d_xu, zg, st));
        CK(cudaEventRecord(e2, st));
        CK((cudaError_t)mpcgpu_b200_kkt_solve_dz_f32(s, m, N, B, d_G, d_C, d_g, d_c, rho, d_lam, d_dz, 10000, tol, d_it, d_ex, nullptr, (void*)st));
        CK(cudaEventRecord(e3, st));
        CK(cudaMemcpyAsync(h_dz, d_dz, B * zg * 4, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(h_it, d_it, B * 4, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(h_ex, d_ex, B, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        if (r >= 0) {
            float a, b2, c2;
            CK(cudaEventElapsedTime(&a, e0, e1)); CK(cudaEventElapsedTime(&b2, e1, e2)); CK(cudaEventElapsedTime(&c2, e2, e3));
            ms_h2d += a / steps; ms_kkt += b2 / steps; ms_solve += c2 / steps;
        }
    }
    const double wall = now_s() - t_begin;
    double it_sum = 0; size_t exits = 0, nonfinite = 0;
    for (uint32_t b = 0; b < B; b++) { it_sum += h_it[b]; exits += h_ex[b]; }
    for (size_t i = 0; i < B * zg; i++) nonfinite += !std::isfinite(h_dz[i]);
    printf("{\"tool\": \"e2e_producer\", \"device\": %d, \"knot_points\": %u, \"batch\": %u, \"exit_tol\": %g, \"steps\": %d, "
           "\"solves_per_s\": %.0f, \"ms_per_step\": %.4f, \"h2d_bytes_per_step\": %zu, \"d2h_bytes_per_step\": %zu, "
           "\"h2d_ms\": %.4f, \"kkt_producer_ms\": %.4f, \"kkt_solve_dz_ms\": %.4f, \"mean_pcg_iters\": %.2f, "
           "\"max_iter_exits\": %zu, \"nonfinite_dz\": %zu, \"gpu_launches_per_step\": %.1f}\n",
           device, N, B, tol, steps, B * steps / wall, wall / steps * 1e3, (B * (zg + zee + s)) * 4, B * (zg * 4 + 5),
           ms_h2d, ms_kkt, ms_solve, it_sum / B, exits, nonfinite, double(mpcgpu_b200_launch_count() - l0) / (steps + warmup));
    return (exits || nonfinite) ? 1 : 0;
}
