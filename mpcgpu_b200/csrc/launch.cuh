// launch.cuh -- one cluster launch of a schur / pcg / fused kernel instantiation (shared by capi.cu and
// capi_f64.cu): opts the kernel in to its shared-memory size (and to cluster sizes > 8) once per device,
// then launches batch * CL CTAs as clusters of CL.
#pragma once
#include <cstdlib>
#include <cstring>

#include "capi_internal.h"
#include "kernels.cuh"

namespace mpcb200 {

// K_FUSED_V2: the fused kernel with PCG variant 2 (S p by recurrence, hidden r.rt reduction) -- an opt-in of the bench
// tiling only (MPCGPU_B200_PCG_VARIANT=2 in the environment; see kkt_device.cuh and DESIGN.md 3 for what it costs in accuracy)
enum Kind { K_SCHUR = 0, K_PCG = 1, K_FUSED = 2, K_FUSED_V2 = 3 };

template <class CF, int KIND>
cudaError_t launch(const BatchArgsT<typename CF::T>& a, cudaStream_t st) {
    auto kern = (KIND == K_SCHUR) ? schur_kernel<CF> : (KIND == K_PCG) ? pcg_kernel<CF> : (KIND == K_FUSED_V2) ? fused_kernel<CF, 2> : fused_kernel<CF>;
    static thread_local int configured_dev = -1;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (configured_dev != dev) {
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CF::SMEM_BYTES);
        if (e != cudaSuccess) return e;
        if (CF::CL > 8) {
            e = cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
            if (e != cudaSuccess) return e;
        }
        configured_dev = dev;
    }
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)a.batch * CF::CL, 1, 1);
    cfg.blockDim = dim3(CF::NT, 1, 1);
    cfg.dynamicSmemBytes = CF::SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CF::CL;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    BatchArgsT<typename CF::T> args = a;
    if ((KIND == K_FUSED || KIND == K_FUSED_V2) && CF::CL == 1) {
        // batches of more than one wave: instance b prefetches the inputs of instance b + (CTAs resident at once) into L2
        static thread_local int resident = 0;
        static const bool on = [] { const char* v = getenv("MPCGPU_B200_L2_PREFETCH"); return !(v && v[0] == '0'); }();
        if (!resident) {
            int sms = 0, per_sm = 0;
            if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess &&
                cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, CF::NT, CF::SMEM_BYTES) == cudaSuccess)
                resident = sms * (per_sm > 0 ? per_sm : 1);
            else { cudaGetLastError(); resident = 148; }
        }
        if (on && a.batch > resident) args.pf_dist = resident;
    }
    e = cudaLaunchKernelEx(&cfg, kern, args);
    if (e == cudaSuccess) detail::count_launch();
    return e;
}

// The cluster "latency tilings" (16 / 32 knots per CTA, 8 lanes per knot) are instantiated in capi_lat.cu
// so that the library's two big translation units compile side by side.
#define MPCB200_LAT_CFGS(X) X(16, 1, 8) X(16, 2, 8) X(16, 4, 8) X(16, 8, 8) X(16, 16, 8) X(32, 16, 8)
#define MPCB200_LAT_DECL(K, C, L)                                                                              \
    extern template cudaError_t launch<Cfg<K, C, L>, K_SCHUR>(const BatchArgsT<float>&, cudaStream_t);          \
    extern template cudaError_t launch<Cfg<K, C, L>, K_PCG>(const BatchArgsT<float>&, cudaStream_t);            \
    extern template cudaError_t launch<Cfg<K, C, L>, K_FUSED>(const BatchArgsT<float>&, cudaStream_t);
#define MPCB200_LAT_DEF(K, C, L)                                                                               \
    template cudaError_t launch<Cfg<K, C, L>, K_SCHUR>(const BatchArgsT<float>&, cudaStream_t);                 \
    template cudaError_t launch<Cfg<K, C, L>, K_PCG>(const BatchArgsT<float>&, cudaStream_t);                   \
    template cudaError_t launch<Cfg<K, C, L>, K_FUSED>(const BatchArgsT<float>&, cudaStream_t);

// per-instance element counts of the reference buffers (include/pcg/sqp.cuh:54-58,111,130)
template <class T>
BatchArgsT<T> make_batch_args(uint32_t s, uint32_t c, uint32_t N, uint32_t batch) {
    BatchArgsT<T> a;
    memset(&a, 0, sizeof(a));
    a.N = (int)N;
    a.batch = (int)batch;
    a.sG = (size_t)(s * s + c * c) * N - c * c;
    a.sC = (size_t)(s * s + s * c) * (N - 1);
    a.sg = (size_t)(s + c) * N - c;
    a.sc = (size_t)s * N;
    a.sS = (size_t)3 * s * s * N;
    a.sv = (size_t)s * N;
    return a;
}

}  // namespace mpcb200
