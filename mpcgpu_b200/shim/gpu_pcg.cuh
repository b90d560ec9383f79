// gpu_pcg.cuh -- drop-in for GBD-PCG's include/gpu_pcg.cuh (absent from the reference
// checkout): the symbols the reference's solver slot uses,
//     pcg<T, STATE_SIZE, KNOT_POINTS>   kernel; argument list include/pcg/sqp.cuh:150-165
//     pcgSharedMemSize<T>(s, N)         include/pcg/sqp.cuh:166
//     checkPcgOccupancy<T>(kernel, block, s, N)   examples/track_iiwa_pcg.cu:24
// plus pcg_config<T> (types.cuh) and gpuErrchk (gpuassert.cuh).
//
// The CALLER fixes the launch geometry: cudaLaunchCooperativeKernel(pcg, grid = knot_points,
// block = PCG_NUM_THREADS (128), args, pcgSharedMemSize) at include/pcg/sqp.cuh:275.  Inside
// that geometry this kernel carries a compile-time __cluster_dims__: the first cluster
// (<= 16 CTAs x 128 threads = 16 knots per CTA, 8 lanes per knot) runs the sm_100a
// cluster/DSMEM PCG of csrc/kkt_device.cuh on the whole horizon; the surplus CTAs of the
// caller's grid retire immediately (nothing waits on them: there is no grid-wide barrier).
// T = double (settings.cuh USE_DOUBLES 1) runs the same cluster kernel instantiated for double while
// the caller's N CTAs stay co-resident at one 124 KB CTA per SM (N <= 64).
// Horizons that do not fit one cluster (N > 256, or N without a usable divisor; N > 64 for double)
// fall back to the cooperative one-CTA-per-knot kernel (csrc/refstyle_pcg.cuh).
//
// Stage 1 behind the same unchanged include/pcg/sqp.cuh: this header also provides the explicit specialisations
// form_schur_system<float> / <double> of the reference's launcher template (include/pcg/linsys_setup.cuh:620-633,
// called at include/pcg/sqp.cuh:243), so the call site binds to the sm_100a cluster kernel
// mpcb200::schur_kernel (csrc/kernels.cuh) instead of the reference's cooperative one-CTA-per-knot
// form_S_gamma_Pinv_kernel -- whichever of the two headers is included first.  The reference's kernel stays
// reachable as its own template for any other type, or for every type with
// -DMPCGPU_B200_SHIM_KEEP_REFERENCE_SCHUR (the parity harness compares the two).
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>
#include <type_traits>

#include "gpuassert.cuh"
#include "types.cuh"
#include "utils.cuh"
#include "glass.cuh"
#include "../csrc/kernels.cuh"
#include "../csrc/refstyle_pcg.cuh"

namespace mpcb200_shim {

constexpr uint32_t KPC = 16;   // knots per 128-thread CTA at 8 lanes per knot

// smallest cluster size that covers N knots, divides the caller's grid (= N) and is <= 16
__host__ __device__ constexpr uint32_t cluster_for(uint32_t N) {
    for (uint32_t cl = (N + KPC - 1) / KPC; cl <= 16; cl++)
        if (cl >= 1 && N % cl == 0) return cl;
    return 0;   // no cluster tiling: cooperative fallback
}

template <typename T, uint32_t S, uint32_t N>
struct Use {
    static constexpr uint32_t CL = cluster_for(N);
    static constexpr bool CLUSTER = S == 14 && CL != 0 &&
        (std::is_same<T, float>::value || (std::is_same<T, double>::value && N <= 64));
    static constexpr uint32_t DIM = CLUSTER ? CL : 1;
};

// The caller's cooperative launch needs all N CTAs of its grid co-resident (as N / cluster whole
// clusters), surplus ones included.  The driver schedules a kernel that holds tensor memory one CTA
// per SM (measured on B200, tools/occupancy_probe.cu: 148 CTAs, 15 clusters of 8, 7 of 16), so the
// Pinv rows go to tensor memory only while the grid fits that way (N <= 64); longer horizons keep
// them in shared memory (3 CTAs per SM).
__host__ __device__ constexpr int tmem_for(uint32_t N) { return N <= 64 ? MPCB200_PINV_TMEM : 0; }

template <typename T, uint32_t N>
using Tiling = mpcb200::Cfg<KPC, (cluster_for(N) ? cluster_for(N) : 1), 8, tmem_for(N), T>;

}  // namespace mpcb200_shim

template <typename T, uint32_t STATE_SIZE_, uint32_t KNOT_POINTS_>
__global__ void
__cluster_dims__(mpcb200_shim::Use<T, STATE_SIZE_, KNOT_POINTS_>::DIM, 1, 1)
__launch_bounds__(128)
pcg(T* d_S, T* d_Pinv, T* d_gamma, T* d_lambda, T* d_r, T* d_p, T* d_v_temp, T* d_eta_new_temp,
    uint32_t* d_pcg_iters, bool* d_pcg_exit, uint32_t max_iter, T exit_tol) {
    extern __shared__ float4 pcg_shim_smem4[];
    if constexpr (mpcb200_shim::Use<T, STATE_SIZE_, KNOT_POINTS_>::CLUSTER) {
        using CF = mpcb200_shim::Tiling<T, KNOT_POINTS_>;
        if (blockIdx.x >= CF::CL) return;            // surplus clusters of the caller's grid
        T* smem = reinterpret_cast<T*>(pcg_shim_smem4);
        mpcb200::RowsOf<CF> R;
        mpcb200::load_rows<CF>(smem, R, d_S, d_Pinv, d_gamma, (int)KNOT_POINTS_);
        mpcb200::pcg_rows<CF>(smem, R, d_lambda, (int)KNOT_POINTS_, max_iter, exit_tol, d_pcg_iters,
                              reinterpret_cast<uint8_t*>(d_pcg_exit));
    } else {
        mpcb200_refstyle::pcg_body<T>(STATE_SIZE_, KNOT_POINTS_, d_S, d_Pinv, d_gamma, d_lambda, d_r, d_p,
                                      d_v_temp, d_eta_new_temp, d_pcg_iters, d_pcg_exit, max_iter,
                                      exit_tol, reinterpret_cast<T*>(pcg_shim_smem4));
    }
}

#if defined(STATE_SIZE) && defined(KNOT_POINTS)
#define MPCB200_SHIM_HAVE_DIMS 1
#else
#define MPCB200_SHIM_HAVE_DIMS 0
#endif

// dynamic shared memory of pcg<T, state_size, knot_points>
template <typename T>
size_t pcgSharedMemSize(uint32_t state_size, uint32_t knot_points) {
    size_t bytes = sizeof(T) * mpcb200_refstyle::smem_elems<T>(state_size);
    const bool cluster = state_size == 14 && mpcb200_shim::cluster_for(knot_points) != 0 &&
                         (std::is_same<T, float>::value || (std::is_same<T, double>::value && knot_points <= 64));
    if (cluster) {
        const size_t cl = mpcb200_shim::cluster_for(knot_points);
        // == Cfg<16, cl, 8, ., T>::SMEM_BYTES: knot_points is a run-time value here, and the cluster size only
        // enters through the two partial-sum buffers (NREDP = cl * 4 warps, rounded up to 4)
        using C1 = mpcb200::Cfg<mpcb200_shim::KPC, 1, 8, 0, T>;
        bytes = C1::SMEM_BYTES + sizeof(T) * 2 * (((cl * 4 + 3) & ~size_t(3)) - C1::NREDP);
    }
#if MPCB200_SHIM_HAVE_DIMS
    // the reference launches with more than the 48 KB default and with cluster sizes up to 16:
    // opt in once per process (callers that skip checkPcgOccupancy still work)
    static bool configured = false;
    if (!configured && state_size == STATE_SIZE && knot_points == KNOT_POINTS) {
        void* k = (void*)pcg<T, STATE_SIZE, KNOT_POINTS>;
        gpuErrchk(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
        if (mpcb200_shim::cluster_for(knot_points) > 8)
            gpuErrchk(cudaFuncSetAttribute(k, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        configured = true;
    }
#endif
    return bytes;
}

// host check before the run (examples/track_iiwa_pcg.cu:24): the cooperative launch of
// `knot_points` CTAs must be co-resident; also opts the kernel in to its shared-memory size.
template <typename T>
bool checkPcgOccupancy(void* kernel, unsigned block, uint32_t state_size, uint32_t knot_points) {
    const size_t smem = pcgSharedMemSize<T>(state_size, knot_points);
    gpuErrchk(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (mpcb200_shim::cluster_for(knot_points) > 8)
        gpuErrchk(cudaFuncSetAttribute(kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    int dev = 0, sms = 0, per_sm = 0, coop = 0;
    gpuErrchk(cudaGetDevice(&dev));
    gpuErrchk(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    gpuErrchk(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
    gpuErrchk(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, (int)block, smem));
    if (!coop || (uint32_t)(per_sm * sms) < knot_points) {
        fprintf(stderr, "pcg: %u CTAs of %u threads (%zu B smem) are not co-resident on this device "
                        "(%d SMs x %d CTAs): reduce KNOT_POINTS\n", knot_points, block, smem, sms, per_sm);
        exit(7);
    }
    return true;
}


// ---- stage 1: form_schur_system behind the unchanged call site (include/pcg/sqp.cuh:243) ---------------------
// the reference's launcher template (include/pcg/linsys_setup.cuh:620-633), declared here so that the
// specialisations below are valid whether or not that header has been seen yet
template <typename T>
void form_schur_system(uint32_t state_size, uint32_t control_size, uint32_t knot_points, T* d_G_dense, T* d_C_dense,
                       T* d_g, T* d_c, T* d_S, T* d_Pinv, T* d_gamma, T rho);

namespace mpcb200_shim {

// our own launch (cluster of ceil(N / 16) CTAs, 16 knots x 8 lanes each; N = 512: 32 knots per CTA), any N
template <typename T, uint32_t N>
struct SchurTiling {
    static constexpr bool F32 = std::is_same<T, float>::value;
    static constexpr uint32_t K = (F32 && N > 256) ? 32 : KPC;
    static constexpr uint32_t CL = (N + K - 1) / K;
    static constexpr bool OK = CL >= 1 && CL <= 16 && (F32 || std::is_same<T, double>::value);
    using CF = mpcb200::Cfg<(int)K, OK ? (int)CL : 1, 8, 0, T>;
};

// form_schur_system<T> for a compile-time horizon: writes d_S, d_Pinv, d_gamma, d_G_dense becomes G^-1.  Legacy
// default stream, like the reference's launch (include/pcg/linsys_setup.cuh:656).
template <typename T, uint32_t N>
void form_schur_system_n(T* d_G_dense, T* d_C_dense, T* d_g, T* d_c, T* d_S, T* d_Pinv, T* d_gamma, T rho) {
    using TL = SchurTiling<T, N>;
    static_assert(TL::OK, "mpcgpu_b200 shim: form_schur_system supports KNOT_POINTS <= 512 (float) / <= 256 (double)");
    using CF = typename TL::CF;
    auto kern = mpcb200::schur_kernel<CF>;
    static int configured_dev = -1;
    int dev = 0;
    gpuErrchk(cudaGetDevice(&dev));
    if (configured_dev != dev) {
        gpuErrchk(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CF::SMEM_BYTES));
        if (CF::CL > 8) gpuErrchk(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        configured_dev = dev;
    }
    constexpr uint32_t s = 14, c = 7;
    mpcb200::BatchArgsT<T> a;
    memset(&a, 0, sizeof(a));
    a.N = (int)N; a.batch = 1;
    a.sG = (size_t)(s * s + c * c) * N - c * c; a.sC = (size_t)(s * s + s * c) * (N - 1);
    a.sg = (size_t)(s + c) * N - c; a.sc = (size_t)s * N; a.sS = (size_t)3 * s * s * N; a.sv = (size_t)s * N;
    a.G = d_G_dense; a.C = d_C_dense; a.g = d_g; a.c = d_c; a.rho = rho;
    a.S = d_S; a.Pinv = d_Pinv; a.gamma = d_gamma;
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(CF::CL, 1, 1);
    cfg.blockDim = dim3(CF::NT, 1, 1);
    cfg.dynamicSmemBytes = CF::SMEM_BYTES;
    cfg.stream = 0;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CF::CL; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    gpuErrchk(cudaLaunchKernelEx(&cfg, kern, a));
}

#if MPCB200_SHIM_HAVE_DIMS
// the run-time signature of the reference, for the horizon this translation unit was compiled for
template <typename T>
void form_schur_system(uint32_t state_size, uint32_t control_size, uint32_t knot_points, T* d_G_dense, T* d_C_dense,
                       T* d_g, T* d_c, T* d_S, T* d_Pinv, T* d_gamma, T rho) {
    if (state_size != 14 || control_size != 7 || knot_points != KNOT_POINTS) {   // loud, never a silent fallback
        fprintf(stderr, "mpcgpu_b200 shim: form_schur_system compiled for state 14 / control 7 / KNOT_POINTS %d, called with "
                        "%u / %u / %u (build with -DMPCGPU_B200_SHIM_KEEP_REFERENCE_SCHUR for the reference's kernel)\n",
                (int)KNOT_POINTS, state_size, control_size, knot_points);
        exit(7);
    }
    form_schur_system_n<T, KNOT_POINTS>(d_G_dense, d_C_dense, d_g, d_c, d_S, d_Pinv, d_gamma, rho);
}
#endif

}  // namespace mpcb200_shim

#if MPCB200_SHIM_HAVE_DIMS && !defined(MPCGPU_B200_SHIM_KEEP_REFERENCE_SCHUR)
#if KNOT_POINTS <= 512
template <>
inline void form_schur_system<float>(uint32_t state_size, uint32_t control_size, uint32_t knot_points, float* d_G_dense,
                                     float* d_C_dense, float* d_g, float* d_c, float* d_S, float* d_Pinv, float* d_gamma,
                                     float rho) {
    mpcb200_shim::form_schur_system<float>(state_size, control_size, knot_points, d_G_dense, d_C_dense, d_g, d_c, d_S, d_Pinv,
                                           d_gamma, rho);
}
#endif
#if KNOT_POINTS <= 256
template <>
inline void form_schur_system<double>(uint32_t state_size, uint32_t control_size, uint32_t knot_points, double* d_G_dense,
                                      double* d_C_dense, double* d_g, double* d_c, double* d_S, double* d_Pinv,
                                      double* d_gamma, double rho) {
    mpcb200_shim::form_schur_system<double>(state_size, control_size, knot_points, d_G_dense, d_C_dense, d_g, d_c, d_S, d_Pinv,
                                            d_gamma, rho);
}
#endif
#endif
