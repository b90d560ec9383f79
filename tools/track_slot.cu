// track_slot.cu -- drives the reference's UNCHANGED closed-loop simulator (include/mpcsim.cuh
// simulateMPC: IIWA end-effector tracking, 2 ms SQP budget per control step) over one of two
// linear-solver slots and prints one JSON line with the tracking statistics:
//
//   -DSLOT_B200=0  the reference's own PCG slot include/pcg/sqp.cuh (sqpSolvePcg) on the L1 shim
//                  headers (its form_schur_system kernel + the shim's pcg<T,S,N>)
//   -DSLOT_B200=1  mpcgpu_b200/shim/b200/sqp.cuh (sqpSolveB200): fused kkt_solve launch, persistent
//                  workspace
//
// mpcsim.cuh picks the slot by name at compile time (include/mpcsim.cuh:21-25,249-277,303-307) and a
// maintainer would add an `#elif LINSYS_SOLVE == 2` there.  Without touching the
// reference we get the same binding by including both slot headers first and renaming the call.
// Built by oracle/build.py where /root/reference exists (needs the reference's headers); usage:
//   track_slot <tol> [start goal]      (cwd must hold examples/trajfiles/ and tmp/results/)
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
// the prelude of include/mpcsim.cuh:16-19, then the slot(s)
#include "integrator.cuh"
#include "settings.cuh"
#include "utils/experiment.cuh"
#include "gpuassert.cuh"
#include "gpu_pcg.cuh"
#include "pcg/sqp.cuh"
#if SLOT_B200
#define MPCGPU_B200_SLOT_COMMON_INCLUDED 1   // common/kkt.cuh, dz.cuh have no include guard; pcg/sqp.cuh brought them
#include "b200/sqp.cuh"
#define sqpSolvePcg sqpSolveB200
#endif
#include "mpcsim.cuh"
#if SLOT_B200
#undef sqpSolvePcg
#endif
#include "dynamics/rbd_plant.cuh"
#include "utils/experiment.cuh"

template <typename V>
static std::vector<linsys_t> flatten(const V& rows) {
    std::vector<linsys_t> out;
    for (const auto& r : rows) out.insert(out.end(), r.begin(), r.end());
    return out;
}

int main(int argc, char** argv) {
    constexpr uint32_t state_size = grid::NUM_JOINTS * 2, control_size = grid::NUM_JOINTS, knot_points = KNOT_POINTS;
    const linsys_t timestep = .015625;
    const float tol = argc > 1 ? float(atof(argv[1])) : 1e-5f;
    const int start = argc > 3 ? atoi(argv[2]) : 0, goal = argc > 3 ? atoi(argv[3]) : 0;

    char name[128];
    snprintf(name, sizeof name, "examples/trajfiles/%d_%d_eepos.traj", start, goal);
    auto ee2d = readCSVToVecVec<linsys_t>(name);
    snprintf(name, sizeof name, "examples/trajfiles/%d_%d_traj.csv", start, goal);
    auto xu2d = readCSVToVecVec<linsys_t>(name);
    if (ee2d.size() < knot_points || xu2d.empty()) { fprintf(stderr, "trajectory files missing or too short\n"); return 2; }
    std::vector<linsys_t> ee = flatten(ee2d), xu = flatten(xu2d);

    linsys_t *d_ee, *d_xu, *d_xs;
    gpuErrchk(cudaMalloc(&d_ee, ee.size() * sizeof(linsys_t)));
    gpuErrchk(cudaMalloc(&d_xu, xu.size() * sizeof(linsys_t)));
    gpuErrchk(cudaMalloc(&d_xs, state_size * sizeof(linsys_t)));
    gpuErrchk(cudaMemcpy(d_ee, ee.data(), ee.size() * sizeof(linsys_t), cudaMemcpyHostToDevice));
    gpuErrchk(cudaMemcpy(d_xu, xu.data(), xu.size() * sizeof(linsys_t), cudaMemcpyHostToDevice));
    gpuErrchk(cudaMemcpy(d_xs, xu.data(), state_size * sizeof(linsys_t), cudaMemcpyHostToDevice));

    std::string prefix = std::string("tmp/results/") + (SLOT_B200 ? "b200_" : "pcg_") + std::to_string(knot_points);
    timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    auto stats = simulateMPC<linsys_t, toplevel_return_type>(state_size, control_size, knot_points, uint32_t(ee2d.size()),
                                                             timestep, d_ee, d_xu, d_xs, start, goal, 0, tol, prefix);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    const auto& per_step = std::get<0>(stats);      // SQP iterations per control step (TIME_LINSYS == 0)
    const auto& errs = std::get<1>(stats);
    auto mean = [](const auto& v) { double a = 0; for (auto x : v) a += double(x); return v.empty() ? 0.0 : a / double(v.size()); };
    double emax = 0; for (auto e : errs) emax = std::max(emax, double(e));
    printf("{\"slot\": \"%s\", \"knot_points\": %u, \"pcg_exit_tol\": %g, \"control_steps\": %zu, "
           "\"%s_mean\": %.4f, \"tracking_err_mean\": %.6f, \"tracking_err_max\": %.6f, "
           "\"final_tracking_err\": %.6f, \"wall_s\": %.2f}\n",
           SLOT_B200 ? "b200/sqp.cuh sqpSolveB200" : "pcg/sqp.cuh sqpSolvePcg (L1 shim)", knot_points, tol, per_step.size(),
           TIME_LINSYS ? "linsys_us" : "sqp_iters", mean(per_step), mean(errs), emax, double(std::get<2>(stats)),
           double(t1.tv_sec - t0.tv_sec) + 1e-9 * double(t1.tv_nsec - t0.tv_nsec));
    cudaFree(d_ee); cudaFree(d_xu); cudaFree(d_xs);
    return 0;
}
