// sqp_batched_check.cu -- K SQP iterations of B perturbed IIWA MPC instances with no host in the loop
// (mpcgpu_b200/shim/b200/sqp_batched.cuh: batched linearisation, fused solve with per-instance rho, the
// 8-step-length line search of all instances in one launch, device-side step selection) against the
// reference's own per-instance sequence of include/pcg/sqp.cuh:203-420 built from ITS kernels:
//   generate_kkt_submatrices -> [our single-instance kkt_solve_dz] -> 8 x cooperative ls_gato_compute_merit
//   -> the host decision / rho adaptation -> xu += alpha dz
// After K iterations xu, rho and the merit of every checked instance must agree bit for bit.
// Prints one JSON line; exit code 1 on mismatch.   sqp_batched_check [N=32] [B=128] [K=4] [tol=1e-4]
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
#include <algorithm>
#include <cooperative_groups.h>

#include "settings.cuh"
#include "gpu_pcg.cuh"
#include "common/kkt.cuh"
#include "utils/experiment.cuh"
#define MPCGPU_B200_SLOT_COMMON_INCLUDED 1
#include "b200/sqp_batched.cuh"

#define CK(x) gpuErrchk(x)

template <typename T>
static T* dalloc(size_t n) { T* d; CK(cudaMalloc(&d, std::max<size_t>(n, 1) * sizeof(T))); return d; }
template <typename T>
static std::vector<T> to_host(const T* d, size_t n) { std::vector<T> h(n); CK(cudaMemcpy(h.data(), d, n * sizeof(T), cudaMemcpyDeviceToHost)); return h; }

__global__ void axpy_step(uint32_t n, float alpha, const float* dz, float* xu) {
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) xu[i] += alpha * dz[i];
}

int main(int argc, char** argv) {
    const uint32_t s = 14, m = 7;
    uint32_t N = argc > 1 ? (uint32_t)atoi(argv[1]) : 32;
    uint32_t B = argc > 2 ? (uint32_t)atoi(argv[2]) : 128;
    const uint32_t K = argc > 3 ? (uint32_t)atoi(argv[3]) : 4;
    float tol = argc > 4 ? (float)atof(argv[4]) : 1e-4f;
    setenv("MPCGPU_B200_LATENCY_TILING", "0", 1);   // batch of B and batch of 1 on the same tiling (same summation order)
    float rho0 = 1e-3f, dt = 0.015625f, mu = 10.0f;   // include/mpcsim.cuh:256, examples/track_iiwa_pcg.cu:19, include/pcg/sqp.cuh:62
    auto ee = readCSVToVecVec<float>("examples/trajfiles/0_0_eepos.traj");
    auto xu = readCSVToVecVec<float>("examples/trajfiles/0_0_traj.csv");
    if (ee.size() < N + 8 || xu.size() < N + 8) { fprintf(stderr, "trajectory too short\n"); return 3; }
    const size_t zG = (size_t)(s * s + m * m) * N - m * m, zC = (size_t)(s * s + s * m) * (N - 1),
                 nz = (size_t)(s + m) * N - m, zc = (size_t)s * N, zv = (size_t)s * N, zee = 6 * (size_t)N;
    std::vector<float> h_xu(B * nz), h_ee(B * zee), h_xs(B * (size_t)s);
    for (uint32_t b = 0; b < B; b++) {
        const uint32_t row0 = b % 8;
        const float amp = 0.002f + 0.02f * (float)b / (float)B;
        size_t o = 0;
        for (uint32_t k = 0; k < N; k++) {
            for (int i = 0; i < 6; i++) h_ee[b * zee + 6 * k + i] = ee[row0 + k][i];
            const size_t take = (k + 1 < N) ? s + m : s;
            for (size_t i = 0; i < take; i++, o++)
                h_xu[b * nz + o] = xu[row0 + k][i] + amp * std::sin(0.37f * (float)o + 1.3f + 0.11f * (float)b);
        }
        for (uint32_t i = 0; i < s; i++) h_xs[b * s + i] = h_xu[b * nz + i] + 0.5f * amp * std::cos(0.9f * (float)i + (float)b);
    }
    float *d_xu = dalloc<float>(B * nz), *d_ee = dalloc<float>(B * zee), *d_xs = dalloc<float>(B * (size_t)s), *d_lam = dalloc<float>(B * zv);
    CK(cudaMemcpy(d_xu, h_xu.data(), h_xu.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_ee, h_ee.data(), h_ee.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_xs, h_xs.data(), h_xs.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemset(d_lam, 0, B * zv * 4));
    void* d_dyn = gato_plant::initializeDynamicsConstMem<float>();

    // ---- batched: K iterations, one stream, no host synchronisation in between -----------------
    b200slot::BatchedSqp<float> sq;
    CK(sq.create(s, m, N, B, rho0));
    cudaStream_t st;
    CK(cudaStreamCreate(&st));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    CK(sq.init_merit(d_xs, d_xu, d_ee, zee, d_dyn, dt, mu, st));
    CK(cudaStreamSynchronize(st));
    std::vector<float> merit0 = to_host(sq.st.merit, B);
    std::vector<double> it_mean(K, 0.0);
    CK(cudaEventRecord(e0, st));
    for (uint32_t k = 0; k < K; k++) CK(sq.iterate(d_xu, d_xs, d_ee, zee, d_lam, d_dyn, dt, mu, 10000, tol, rho0, st));
    CK(cudaEventRecord(e1, st));
    CK(cudaStreamSynchronize(st));
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    std::vector<float> b_xu = to_host(d_xu, B * nz), b_rho = to_host(sq.st.rho, B), b_merit = to_host(sq.st.merit, B), b_step = to_host(sq.st.step, B);
    std::vector<uint32_t> b_flags = to_host(sq.st.flags, B);
    std::vector<uint32_t> b_iters = to_host(sq.iters, B);
    size_t accepted = 0, improved = 0, nonfinite = 0, resets = 0;
    double it_sum = 0;
    for (uint32_t b = 0; b < B; b++) {
        accepted += b_step[b] != 0.f; improved += b_merit[b] < merit0[b]; resets += b_flags[b] & 1u; it_sum += b_iters[b];
        for (size_t i = 0; i < nz; i++) nonfinite += !std::isfinite(b_xu[b * nz + i]);
    }

    // ---- the reference's sequence, one instance at a time ---------------------------------------
    float *r_G = dalloc<float>(zG), *r_C = dalloc<float>(zC), *r_g = dalloc<float>(nz), *r_c = dalloc<float>(zc);
    float *r_xu = dalloc<float>(nz), *r_lam = dalloc<float>(zv), *r_dz = dalloc<float>(nz);
    float *r_merits = dalloc<float>(8), *r_tmp = dalloc<float>(8 * (size_t)N);
    uint32_t* r_it = dalloc<uint32_t>(1);
    uint8_t* r_ex = dalloc<uint8_t>(1);
    CK(cudaMemset(r_c, 0, zc * 4));
    void* ls_kernel = (void*)ls_gato_compute_merit<float>;
    const size_t ls_smem = get_merit_smem_size<float>(s, N);          // include/pcg/sqp.cuh:343
    const size_t kkt_smem = 2 * get_kkt_smem_size<float>(s, m);
    const uint32_t ncheck = std::min<uint32_t>(B, 16);
    size_t mism_xu = 0, mism_rho = 0, mism_merit = 0;
    uint32_t ks = s, kc = m, kn = N;
    for (uint32_t i = 0; i < ncheck; i++) {
        const uint32_t b = (uint32_t)((uint64_t)i * B / ncheck);
        if (b_flags[b] & 1u) continue;                 // the reference leaves its loop when rho is reset
        CK(cudaMemcpy(r_xu, h_xu.data() + b * nz, nz * 4, cudaMemcpyHostToDevice));
        CK(cudaMemset(r_lam, 0, zv * 4));
        float rho = rho0, drho = 1.f, merit_now = merit0[b];
        const float rho_factor = RHO_FACTOR, rho_min = RHO_MIN, rho_max = RHO_MAX;
        float* xs_b = d_xs + b * s; float* ee_b = d_ee + b * zee;
        for (uint32_t k = 0; k < K; k++) {
            generate_kkt_submatrices<float><<<N, KKT_THREADS, kkt_smem>>>(s, m, N, r_G, r_C, r_g, r_c, d_dyn, dt, ee_b, xs_b, r_xu);
            CK(cudaPeekAtLastError());
            CK((cudaError_t)mpcgpu_b200_kkt_solve_dz_f32(s, m, N, 1, r_G, r_C, r_g, r_c, rho, r_lam, r_dz, 10000, tol, r_it, r_ex, nullptr, nullptr));
            for (uint32_t p = 0; p < 8; p++) {
                void* args[] = {&ks, &kc, &kn, &xs_b, &r_xu, &ee_b, &mu, &dt, &d_dyn, &r_dz, &p, &r_merits, &r_tmp};
                CK(cudaLaunchCooperativeKernel(ls_kernel, N, MERIT_THREADS, args, ls_smem, 0));
            }
            std::vector<float> hm = to_host(r_merits, 8);
            uint32_t best = 0;
            float best_merit = merit_now;
            for (uint32_t p = 0; p < 8; p++) if (hm[p] < best_merit) { best_merit = hm[p]; best = p; }
            if (best_merit == merit_now) {
                drho = std::max(drho * rho_factor, rho_factor);
                rho = std::max(rho * drho, rho_min);
                if (rho > rho_max) break;
                continue;
            }
            drho = std::min(drho / rho_factor, 1 / rho_factor);
            rho = std::max(rho * drho, rho_min);
            axpy_step<<<1, 256>>>((uint32_t)nz, -1.0f / (float)(1u << best), r_dz, r_xu);
            merit_now = best_merit;
        }
        CK(cudaDeviceSynchronize());
        auto a = to_host(r_xu, nz);
        mism_xu += memcmp(a.data(), b_xu.data() + b * nz, nz * 4) != 0;
        mism_rho += memcmp(&rho, &b_rho[b], 4) != 0;
        mism_merit += memcmp(&merit_now, &b_merit[b], 4) != 0;
    }
    printf("{\"tool\": \"sqp_batched_check\", \"knot_points\": %u, \"batch\": %u, \"sqp_iterations\": %u, \"exit_tol\": %g, "
           "\"instances_checked\": %u, \"xu_mismatches_vs_reference_sequence\": %zu, \"rho_mismatches\": %zu, \"merit_mismatches\": %zu, "
           "\"nonfinite_xu\": %zu, \"instances_with_step_in_last_iteration\": %zu, \"instances_with_lower_merit\": %zu, \"rho_resets\": %zu, "
           "\"mean_pcg_iters_last\": %.2f, \"batched_ms_per_sqp_iteration\": %.4f, \"instance_sqp_iterations_per_s\": %.0f}\n",
           N, B, K, tol, ncheck, mism_xu, mism_rho, mism_merit, nonfinite, accepted, improved, resets, it_sum / B, ms / K,
           (double)B * K / (ms * 1e-3));
    return (mism_xu || mism_rho || mism_merit || nonfinite) ? 1 : 0;
}
