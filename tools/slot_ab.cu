// slot_ab.cu -- TEST TOOL: one SQP iteration of the reference's PCG slot against one SQP iteration of the
// b200 slot on IDENTICAL (xu, lambda0, rho), chained along a real IIWA tracking problem.
//
//   A  include/pcg/sqp.cuh sqpSolvePcg   -- the reference's UNCHANGED slot: its generate_kkt_submatrices, its
//      form_schur_system kernel, pcg<T,14,N> of the L1 shim launched exactly as :275, its compute_dz, its 8
//      line-search kernels, its host decision and cublasSaxpy
//   B  mpcgpu_b200/shim/b200/sqp.cuh sqpSolveB200 -- fused kkt_solve_dz launch, CUDA-graphed iteration
//
// Both are compiled with SQP_MAX_ITER redefined to 1 (a macro override in THIS translation unit; the reference
// files are untouched) and an unbounded time budget, so that every call is exactly one SQP iteration.  For
// each window of the 0_0 trajectory: K chained iterations; before each one both slots get copies of the same
// state; the state then continues from A's result (so lambda0 is warm, rho adapts, as in the closed loop).
// One JSON line: per-iteration PCG iteration counts, relative differences of lambda and of the applied step
// xu_new - xu_old (= alpha dz), equality of the accepted step length and of rho.
//
//   slot_ab <tol> <K> [perturb]        (cwd must hold examples/trajfiles/)
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>
#include <numeric>
#include <algorithm>
#include <cmath>
#include <iomanip>
#include <fstream>
#include <iostream>
#include <cstdint>
#include <cublas_v2.h>
#include <cuda_runtime.h>
#include <tuple>
#include <time.h>
#include "integrator.cuh"
#include "settings.cuh"
#undef SQP_MAX_ITER
#define SQP_MAX_ITER 1
#include "utils/experiment.cuh"
#include "gpuassert.cuh"
#ifndef SLOT_AB_L1_SCHUR     // A's stage 1: the reference's own form_schur_system kernel (default) or the L1 shim's
#define MPCGPU_B200_SHIM_KEEP_REFERENCE_SCHUR 1
#endif
#include "gpu_pcg.cuh"
#include "pcg/sqp.cuh"
#define MPCGPU_B200_SLOT_COMMON_INCLUDED 1
#include "b200/sqp.cuh"

template <typename V>
static std::vector<float> flatten(const V& rows) {
    std::vector<float> out;
    for (const auto& r : rows) out.insert(out.end(), r.begin(), r.end());
    return out;
}
static double rel(const std::vector<float>& a, const std::vector<float>& b) {
    double n = 0, d = 0;
    for (size_t i = 0; i < a.size(); i++) { d += double(a[i] - b[i]) * double(a[i] - b[i]); n += double(b[i]) * double(b[i]); }
    return n > 0 ? std::sqrt(d / n) : std::sqrt(d);
}
static std::vector<float> down(const float* d, size_t n) {
    std::vector<float> h(n);
    gpuErrchk(cudaMemcpy(h.data(), d, n * sizeof(float), cudaMemcpyDeviceToHost));
    return h;
}

int main(int argc, char** argv) {
    static_assert(std::is_same<linsys_t, float>::value, "slot_ab compares the fp32 slots");
    constexpr uint32_t s = grid::NUM_JOINTS * 2, c = grid::NUM_JOINTS, N = KNOT_POINTS;
    const float timestep = .015625f;
    const float tol = argc > 1 ? float(atof(argv[1])) : 1e-4f;
    const int K = argc > 2 ? atoi(argv[2]) : 6;
    const float perturb = argc > 3 ? float(atof(argv[3])) : 0.02f;
    auto ee2d = readCSVToVecVec<float>("examples/trajfiles/0_0_eepos.traj");
    auto xu2d = readCSVToVecVec<float>("examples/trajfiles/0_0_traj.csv");
    if (ee2d.size() < N + 400 || xu2d.size() < N + 400) { fprintf(stderr, "trajectory too short\n"); return 2; }
    const uint32_t nz = (s + c) * N - c, nl = s * N;
    float *d_ee, *xu0, *xuA, *xuB, *l0, *lA, *lB;
    gpuErrchk(cudaMalloc(&d_ee, 6 * N * 4));
    for (float** p : {&xu0, &xuA, &xuB}) gpuErrchk(cudaMalloc(p, nz * 4));
    for (float** p : {&l0, &lA, &lB}) gpuErrchk(cudaMalloc(p, nl * 4));
    void* dyn = gato_plant::initializeDynamicsConstMem<float>();
    pcg_config<float> cfg;
    cfg.pcg_block = PCG_NUM_THREADS; cfg.pcg_exit_tol = tol; cfg.pcg_max_iter = PCG_MAX_ITER;

    double worst_lam = 0, worst_step = 0, worst_xu = 0;
    int n_iter = 0, n_iters_pm1 = 0, n_same_rho = 0, n_same_accept = 0, max_iter_diff = 0, n_accepted = 0, n_same_len = 0;
    std::string per = "[";
    for (uint32_t row0 : {0u, 100u, 200u, 300u}) {
        std::vector<float> h_ee, h_xu;
        for (uint32_t k = 0; k < N; k++) {
            h_ee.insert(h_ee.end(), ee2d[row0 + k].begin(), ee2d[row0 + k].begin() + 6);
            const size_t take = (k + 1 < N) ? s + c : s;
            h_xu.insert(h_xu.end(), xu2d[row0 + k].begin(), xu2d[row0 + k].begin() + take);
        }
        for (size_t i = s; i < h_xu.size(); i++) h_xu[i] += perturb * std::sin(0.37f * (float)i + 1.3f);   // an iterate off the nominal trajectory
        gpuErrchk(cudaMemcpy(d_ee, h_ee.data(), h_ee.size() * 4, cudaMemcpyHostToDevice));
        gpuErrchk(cudaMemcpy(xu0, h_xu.data(), nz * 4, cudaMemcpyHostToDevice));
        gpuErrchk(cudaMemset(l0, 0, nl * 4));
        float rho = 1e-3f;
        for (int it = 0; it < K; it++) {
            gpuErrchk(cudaMemcpy(xuA, xu0, nz * 4, cudaMemcpyDeviceToDevice)); gpuErrchk(cudaMemcpy(xuB, xu0, nz * 4, cudaMemcpyDeviceToDevice));
            gpuErrchk(cudaMemcpy(lA, l0, nl * 4, cudaMemcpyDeviceToDevice));   gpuErrchk(cudaMemcpy(lB, l0, nl * 4, cudaMemcpyDeviceToDevice));
            float rhoA = rho, rhoB = rho;
            auto ra = sqpSolvePcg<float>(s, c, N, timestep, d_ee, lA, xuA, dyn, cfg, rhoA, 1e-3f);
            auto rb = sqpSolveB200<float>(s, c, N, timestep, d_ee, lB, xuB, dyn, cfg, rhoB, 1e-3f);
            const auto h0 = down(xu0, nz), hA = down(xuA, nz), hB = down(xuB, nz), hlA = down(lA, nl), hlB = down(lB, nl);
            std::vector<float> stA(nz), stB(nz);
            double nA = 0, nB = 0;
            for (uint32_t i = 0; i < nz; i++) { stA[i] = hA[i] - h0[i]; stB[i] = hB[i] - h0[i]; nA += double(stA[i]) * stA[i]; nB += double(stB[i]) * stB[i]; }
            const int ia = std::get<0>(ra).empty() ? -1 : std::get<0>(ra)[0], ib = std::get<0>(rb).empty() ? -1 : std::get<0>(rb)[0];
            const bool accA = nA > 0, accB = nB > 0;
            const double dl = rel(hlB, hlA), dst = (accA && accB) ? rel(stB, stA) : 0.0, dxu = rel(hB, hA);
            n_iter++;
            n_iters_pm1 += std::abs(ia - ib) <= 1; max_iter_diff = std::max(max_iter_diff, std::abs(ia - ib));
            n_same_rho += rhoA == rhoB; n_same_accept += accA == accB; n_accepted += accA;
            // the step length alpha is one of 1, 1/2, ..., 1/128: the same one iff the two steps have (nearly) the same norm
            const bool same_len = accA && accB && std::fabs(std::sqrt(nB / nA) - 1.0) < 0.2;
            n_same_len += same_len;
            worst_lam = std::max(worst_lam, dl); worst_step = std::max(worst_step, dst); worst_xu = std::max(worst_xu, dxu);
            char buf[256];
            snprintf(buf, sizeof buf, "%s{\"row0\": %u, \"it\": %d, \"pcg_iters\": [%d, %d], \"rho\": [%g, %g], \"accepted\": [%d, %d], "
                     "\"lambda_rel\": %.3g, \"step_rel\": %.3g}", per.size() > 1 ? ", " : "", row0, it, ia, ib, rhoA, rhoB, int(accA), int(accB), dl, dst);
            per += buf;
            gpuErrchk(cudaMemcpy(xu0, xuA, nz * 4, cudaMemcpyDeviceToDevice));
            gpuErrchk(cudaMemcpy(l0, lA, nl * 4, cudaMemcpyDeviceToDevice));
            rho = rhoA;
        }
    }
    per += "]";
    printf("{\"knot_points\": %u, \"pcg_exit_tol\": %g, \"sqp_iterations_compared\": %d, \"pcg_iters_within_1\": %d, "
           "\"max_pcg_iter_diff\": %d, \"same_rho\": %d, \"same_accept_decision\": %d, \"accepted_steps\": %d, \"same_step_length\": %d, "
           "\"lambda_rel_max\": %.4g, \"step_rel_max\": %.4g, \"xu_rel_max\": %.4g, \"per_iteration\": %s}\n",
           N, tol, n_iter, n_iters_pm1, max_iter_diff, n_same_rho, n_same_accept, n_accepted, n_same_len, worst_lam, worst_step, worst_xu, per.c_str());
    return 0;
}
