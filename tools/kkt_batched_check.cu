// kkt_batched_check.cu -- BASELINE.json configs[4] end to end on real IIWA dynamics:
//   B MPC instances with perturbed start states / iterates
//     -> b200slot::generate_kkt_submatrices_batched   (mpcgpu_b200/shim/b200/kkt_batched.cuh, ONE launch)
//     -> mpcgpu_b200_kkt_solve_dz_f32(batch = B)       (ONE launch: Schur + PCG + compute_dz)
// and the checks that make it a drop-in:
//   (1) every instance's G, C, g, c equal, bit for bit, what the reference's OWN kernel
//       generate_kkt_submatrices (include/common/kkt.cuh:22-163) writes for that instance alone;
//   (2) the batched solve equals, bit for bit, B single-instance solves (lambda, dz, iterations).
// Prints one JSON line; exit code 1 on any mismatch.  Built by oracle/build.py where /root/reference
// exists (it needs the reference's headers); cwd must hold examples/trajfiles/.
//   kkt_batched_check [N=64] [B=256] [tol=1e-4]
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
#include <cooperative_groups.h>

#include "settings.cuh"
#include "gpu_pcg.cuh"
#include "common/kkt.cuh"
#include "utils/experiment.cuh"
#define MPCGPU_B200_SLOT_COMMON_INCLUDED 1
#include "b200/kkt_batched.cuh"
#include "mpcgpu_b200.h"

#define CK(x) gpuErrchk(x)

template <typename T>
static T* dalloc(size_t n) { T* d; CK(cudaMalloc(&d, std::max<size_t>(n, 1) * sizeof(T))); return d; }
template <typename T>
static std::vector<T> to_host(const T* d, size_t n) { std::vector<T> h(n); CK(cudaMemcpy(h.data(), d, n * sizeof(T), cudaMemcpyDeviceToHost)); return h; }

int main(int argc, char** argv) {
    const uint32_t s = 14, m = 7;
    const uint32_t N = argc > 1 ? (uint32_t)atoi(argv[1]) : 64;
    const uint32_t B = argc > 2 ? (uint32_t)atoi(argv[2]) : 256;
    const float tol = argc > 3 ? (float)atof(argv[3]) : 1e-4f;
    // one tiling for every call, so that a batch of B and a batch of 1 run the same arithmetic (the library
    // would otherwise pick its cluster "latency tiling" for the single solves: same math, other summation order)
    setenv("MPCGPU_B200_LATENCY_TILING", "0", 1);
    const float rho = 1e-3f, timestep = 0.015625f;          // include/mpcsim.cuh:256, examples/track_iiwa_pcg.cu:19
    auto ee = readCSVToVecVec<float>("examples/trajfiles/0_0_eepos.traj");
    auto xu = readCSVToVecVec<float>("examples/trajfiles/0_0_traj.csv");
    if (ee.size() < N + 8 || xu.size() < N + 8) { fprintf(stderr, "trajectory too short\n"); return 3; }
    const size_t zG = (size_t)(s * s + m * m) * N - m * m, zC = (size_t)(s * s + s * m) * (N - 1),
                 zg = (size_t)(s + m) * N - m, zc = (size_t)s * N, zv = (size_t)s * N, zee = 6 * (size_t)N;

    // instance b: window of the nominal trajectory starting at row (b % 8), iterate perturbed by a
    // deterministic ripple whose amplitude grows with b, start state off the iterate's x_0
    std::vector<float> h_xu(B * zg), h_ee(B * zee), h_xs(B * (size_t)s);
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
    CK(cudaMemcpy(d_xu, h_xu.data(), h_xu.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_ee, h_ee.data(), h_ee.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_xs, h_xs.data(), h_xs.size() * 4, cudaMemcpyHostToDevice));
    void* d_dyn = gato_plant::initializeDynamicsConstMem<float>();

    float *d_G = dalloc<float>(B * zG), *d_C = dalloc<float>(B * zC), *d_g = dalloc<float>(B * zg), *d_c = dalloc<float>(B * zc);
    CK(cudaMemset(d_G, 0xff, B * zG * 4)); CK(cudaMemset(d_C, 0xff, B * zC * 4));
    CK(cudaMemset(d_g, 0xff, B * zg * 4)); CK(cudaMemset(d_c, 0xff, B * zc * 4));
    cudaEvent_t e0, e1, e2;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1)); CK(cudaEventCreate(&e2));

    // ---- batched producer --------------------------------------------------------------------
    CK(b200slot::launch_kkt_batched<float>(s, m, N, B, d_G, d_C, d_g, d_c, d_dyn, timestep, d_ee, zee, d_xs, d_xu, zg, 0));
    CK(cudaDeviceSynchronize());
    std::vector<float> hG = to_host(d_G, B * zG), hC = to_host(d_C, B * zC), hg = to_host(d_g, B * zg), hc = to_host(d_c, B * zc);

    // ---- (1) the reference's own kernel, one instance at a time ---------------------------------
    float *r_G = dalloc<float>(zG), *r_C = dalloc<float>(zC), *r_g = dalloc<float>(zg), *r_c = dalloc<float>(zc);
    size_t kkt_mismatch = 0;
    const uint32_t ncheck = std::min<uint32_t>(B, 32);
    for (uint32_t i = 0; i < ncheck; i++) {
        const uint32_t b = (uint32_t)((uint64_t)i * B / ncheck);
        generate_kkt_submatrices<float><<<N, KKT_THREADS, 2 * get_kkt_smem_size<float>(s, m)>>>(
            s, m, N, r_G, r_C, r_g, r_c, d_dyn, timestep, d_ee + b * zee, d_xs + b * s, d_xu + b * zg);   // include/pcg/sqp.cuh:212
        CK(cudaPeekAtLastError());
        CK(cudaDeviceSynchronize());
        auto a = to_host(r_G, zG); kkt_mismatch += memcmp(a.data(), hG.data() + b * zG, zG * 4) != 0;
        a = to_host(r_C, zC); kkt_mismatch += memcmp(a.data(), hC.data() + b * zC, zC * 4) != 0;
        a = to_host(r_g, zg); kkt_mismatch += memcmp(a.data(), hg.data() + b * zg, zg * 4) != 0;
        a = to_host(r_c, zc); kkt_mismatch += memcmp(a.data(), hc.data() + b * zc, zc * 4) != 0;
    }

    // ---- batched linearise + solve step, timed (KKT blocks regenerated every repetition: the solve
    // overwrites G with G^-1) ------------------------------------------------------------------------
    float *d_lam = dalloc<float>(B * zv), *d_dz = dalloc<float>(B * zg);
    uint32_t* d_it = dalloc<uint32_t>(B);
    uint8_t* d_ex = dalloc<uint8_t>(B);
    float ms_kkt = 0.f, ms_solve = 0.f;
    const int reps = 10;
    for (int r = -2; r < reps; r++) {
        CK(cudaMemsetAsync(d_lam, 0, B * zv * 4));
        CK(cudaEventRecord(e0));
        CK(b200slot::launch_kkt_batched<float>(s, m, N, B, d_G, d_C, d_g, d_c, d_dyn, timestep, d_ee, zee, d_xs, d_xu, zg, 0));
        CK(cudaEventRecord(e1));
        CK((cudaError_t)mpcgpu_b200_kkt_solve_dz_f32(s, m, N, B, d_G, d_C, d_g, d_c, rho, d_lam, d_dz, 10000, tol, d_it, d_ex, nullptr, nullptr));
        CK(cudaEventRecord(e2));
        CK(cudaDeviceSynchronize());
        float a, b2;
        CK(cudaEventElapsedTime(&a, e0, e1)); CK(cudaEventElapsedTime(&b2, e1, e2));
        if (r >= 0) { ms_kkt += a / reps; ms_solve += b2 / reps; }
    }
    std::vector<float> lam = to_host(d_lam, B * zv), dz = to_host(d_dz, B * zg);
    std::vector<uint32_t> it = to_host(d_it, B);
    std::vector<uint8_t> ex = to_host(d_ex, B);

    // ---- (2) the same instances solved one at a time --------------------------------------------
    size_t solve_mismatch = 0, nonfinite = 0, exits = 0;
    double it_sum = 0;
    uint32_t it_min = ~0u, it_max = 0;
    for (uint32_t b = 0; b < B; b++) {
        it_sum += it[b]; it_min = std::min(it_min, it[b]); it_max = std::max(it_max, it[b]); exits += ex[b];
        for (size_t i = 0; i < zg; i++) nonfinite += !std::isfinite(dz[b * zg + i]);
    }
    float *s_lam = dalloc<float>(zv), *s_dz = dalloc<float>(zg);
    for (uint32_t i = 0; i < ncheck; i++) {
        const uint32_t b = (uint32_t)((uint64_t)i * B / ncheck);
        CK(cudaMemcpy(r_G, hG.data() + b * zG, zG * 4, cudaMemcpyHostToDevice));
        CK(cudaMemset(s_lam, 0, zv * 4));
        CK((cudaError_t)mpcgpu_b200_kkt_solve_dz_f32(s, m, N, 1, r_G, d_C + b * zC, d_g + b * zg, d_c + b * zc, rho, s_lam, s_dz,
                                                      10000, tol, d_it, d_ex, nullptr, nullptr));
        CK(cudaDeviceSynchronize());
        auto a = to_host(s_lam, zv); solve_mismatch += memcmp(a.data(), lam.data() + b * zv, zv * 4) != 0;
        a = to_host(s_dz, zg); solve_mismatch += memcmp(a.data(), dz.data() + b * zg, zg * 4) != 0;
        auto one = to_host(d_it, 1); solve_mismatch += one[0] != it[b];
    }
    printf("{\"tool\": \"kkt_batched_check\", \"knot_points\": %u, \"batch\": %u, \"exit_tol\": %g, \"instances_checked\": %u, "
           "\"kkt_mismatches_vs_reference_kernel\": %zu, \"solve_mismatches_vs_single\": %zu, \"nonfinite_dz\": %zu, "
           "\"max_iter_exits\": %zu, \"pcg_iters\": {\"mean\": %.2f, \"min\": %u, \"max\": %u}, \"kkt_batched_ms\": %.4f, "
           "\"kkt_solve_dz_ms\": %.4f, \"linearise_and_solve_steps_per_s\": %.0f}\n",
           N, B, tol, ncheck, kkt_mismatch, solve_mismatch, nonfinite, exits, it_sum / B, it_min, it_max, ms_kkt, ms_solve,
           B / ((ms_kkt + ms_solve) * 1e-3));
    return (kkt_mismatch || solve_mismatch || nonfinite) ? 1 : 0;
}
