// noalloc.cu -- ONE entry point of the C ABI that allocates nothing, for callers that cannot call cudaMalloc (code
// under CUDA graph capture, tight serving loops, harnesses that forbid it).  No counterpart in the
// reference.
//
//   mpcgpu_b200_pcg_refstyle_noalloc_f32 = mpcgpu_b200_pcg_refstyle_f32 (the reference-STYLE cooperative PCG, the
//   one kernel of this library that needs global scratch: d_r, d_p, d_v_temp, d_eta_new_temp) with that scratch in
//   a file-scope device array instead of a caller-supplied d_ws.
//
// The array is sized for the largest horizon the ABI serves (s = 14, N <= 512: 2*14*512 + 2*512 floats = 61 440 B);
// every GPU that loads the library holds it until exit.  EVERY call on a device shares it and it keeps the last
// call's contents: the kernel writes every scratch element before it reads it (r and p rows before the first
// halo load, the per-knot partial sums before each reduction), so nothing has to be cleared -- but two calls
// that overlap on the same GPU (different streams, or another host thread) race on it.  Serialise them.
// In host code `g_noalloc_scratch` is only a host-side stand-in for the symbol, not a device address: the address
// is resolved once per device with cudaGetSymbolAddress (that device current) and passed to the kernel.
// The entry point calls no cudaMalloc / cudaFree and no allocating library routine.
#include "../../include/mpcgpu_b200.h"

#include <cuda_runtime.h>

#include <mutex>

#include "capi_internal.h"
#include "refstyle_pcg.cuh"

namespace {

constexpr uint32_t kState = 14;
constexpr uint32_t kMaxKnots = 512;
constexpr size_t kScratchFloats = (size_t)2 * kState * kMaxKnots + 2 * kMaxKnots;   // d_r | d_p | d_v_temp | d_eta_new_temp
constexpr int kMaxDevices = 64;

static __device__ __align__(256) float g_noalloc_scratch[kScratchFloats];

__global__ void refstyle_noalloc_kernel(uint32_t s, uint32_t N, const float* d_S, const float* d_Pinv, const float* d_gamma,
                                        float* d_lambda, float* scratch, uint32_t* d_iters, bool* d_exit, uint32_t max_iter,
                                        float tol) {
    extern __shared__ float rs_noalloc_smem[];
    float* d_r = scratch;
    float* d_p = d_r + (size_t)s * N;
    float* d_v = d_p + (size_t)s * N;
    float* d_eta = d_v + N;
    mpcb200_refstyle::pcg_body<float>(s, N, d_S, d_Pinv, d_gamma, d_lambda, d_r, d_p, d_v, d_eta, d_iters, d_exit,
                                      max_iter, tol, rs_noalloc_smem);
}

std::mutex g_mu;
float* g_scratch_addr[kMaxDevices] = {};    // device address of g_noalloc_scratch, per device

// set-up, once per device, with that device current
cudaError_t scratch_for_current_device(float** out) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_addr[dev]) {
        float* p = nullptr;
        e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        g_scratch_addr[dev] = p;
    }
    *out = g_scratch_addr[dev];
    return cudaSuccess;
}

}  // namespace

extern "C" {

uint32_t mpcgpu_b200_noalloc_max_knots(void) { return kMaxKnots; }

int mpcgpu_b200_pcg_refstyle_noalloc_f32(uint32_t s, uint32_t N, const float* d_S, const float* d_Pinv,
                                         const float* d_gamma, float* d_lambda, uint32_t max_iter, float exit_tol,
                                         uint32_t* d_iters, uint8_t* d_max_iter_exit, void* stream) {
    using mpcb200::detail::bad;
    using mpcb200::detail::fail;
    if (s != kState)
        return bad("mpcgpu_b200_pcg_refstyle_noalloc_f32: state_size must be 14 (the static scratch is sized for s = 14)");
    if (N < 1 || N > kMaxKnots)
        return bad("mpcgpu_b200_pcg_refstyle_noalloc_f32: knot_points outside [1, 512]: the static scratch holds at most "
                   "512 knots (61 440 B); use mpcgpu_b200_pcg_refstyle_f32 with a caller-owned d_ws for longer horizons");
    if (!d_S || !d_Pinv || !d_gamma || !d_lambda || !d_iters || !d_max_iter_exit)
        return bad("mpcgpu_b200_pcg_refstyle_noalloc_f32: null pointer");
    float* scratch = nullptr;
    cudaError_t e = scratch_for_current_device(&scratch);
    if (e != cudaSuccess) return fail(e, "mpcgpu_b200_pcg_refstyle_noalloc_f32: cudaGetSymbolAddress");
    bool* d_exit = reinterpret_cast<bool*>(d_max_iter_exit);
    size_t smem = sizeof(float) * mpcb200_refstyle::smem_elems<float>(s);       // 5 464 B: below the 48 KB default
    void* args[] = {&s, &N, &d_S, &d_Pinv, &d_gamma, &d_lambda, &scratch, &d_iters, &d_exit, &max_iter, &exit_tol};
    e = cudaLaunchCooperativeKernel((void*)refstyle_noalloc_kernel, dim3(N), dim3(128), args, smem, (cudaStream_t)stream);
    if (e != cudaSuccess) return fail(e, "mpcgpu_b200_pcg_refstyle_noalloc_f32");
    mpcb200::detail::count_launch();
    return 0;
}

}  // extern "C"
