// capi.cu -- C ABI (include/mpcgpu_b200.h) and kernel dispatch.
// No torch, no CPU fallback: every compute entry point launches sm_100a kernels or fails.
#include "../../include/mpcgpu_b200.h"

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "launch.cuh"
#include "refstyle_pcg.cuh"

#ifndef MPCB200_MINIMAL
namespace mpcb200 {
MPCB200_LAT_CFGS(MPCB200_LAT_DECL)      // defined in capi_lat.cu
}
#endif

using namespace mpcb200;

namespace {

thread_local std::string g_err;
std::atomic<uint64_t> g_launches{0};

int fail(cudaError_t e, const char* what) {
    g_err = std::string(what) + ": " + cudaGetErrorString(e);
    return (int)e;
}
int bad(const char* what) {
    g_err = what;
    return (int)cudaErrorInvalidValue;
}
}  // namespace
namespace mpcb200 { namespace detail {
int fail(cudaError_t e, const char* what) { return ::fail(e, what); }
int bad(const char* what) { return ::bad(what); }
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
} }
namespace {


// (KPC, CL, LPK) tiling: N <= KPC*CL.
//  * throughput tiling (many instances): 64 knots per CTA -- a 64-knot instance lives in ONE SM,
//    no cluster barrier at all; lanes per knot from MPCGPU_B200_LPK (4 or 8).
//  * latency tiling (the whole batch fits one wave of 16-knot CTAs): 16 knots per 128-thread
//    CTA, cluster of N/16 CTAs (<= 16) -- the per-iteration shared-memory time scales with the
//    knots per CTA, so a single solve spreads over as many SMs as a cluster allows.
template <int KIND>
cudaError_t dispatch(const BatchArgs& a, cudaStream_t st) {
    static const int lpk = [] {
        const char* e = getenv("MPCGPU_B200_LPK");
        return (e && atoi(e) == 8) ? 8 : 4;      // measured on B200: 839k solves/s (4) vs 770k (8)
    }();
    static const int lat = [] {
        const char* e = getenv("MPCGPU_B200_LATENCY_TILING");
        return e ? atoi(e) : -1;      // -1: automatic, 0: never, 1: always
    }();
    static const bool var2 = [] {     // opt-in: PCG variant 2 in the fused kernel of the 64-knot throughput tiling
        const char* e = getenv("MPCGPU_B200_PCG_VARIANT");
        return e && atoi(e) == 2;
    }();
    (void)var2;
    const int N = a.N;
#ifdef MPCB200_MINIMAL   // experiment builds (tools/ab_build.sh): the bench tiling only, a minute of nvcc
#ifdef MPCB200_MIN64_K     // the tiling under test for N <= 64 (default: the bench tiling)
    if (N <= 64) return launch<Cfg<MPCB200_MIN64_K, MPCB200_MIN64_C, MPCB200_MIN64_L, MPCB200_MIN64_TM>, KIND>(a, st);
#else
    if (N <= 64) return launch<Cfg<64, 1, 4>, KIND>(a, st);
#endif
#ifdef MPCB200_MIN128_K    // -DMPCB200_MIN128_K=32 -DMPCB200_MIN128_C=4 -DMPCB200_MIN128_L=8: the tiling under test for 64 < N <= 128
    if (N <= 128) return launch<Cfg<MPCB200_MIN128_K, MPCB200_MIN128_C, MPCB200_MIN128_L>, KIND>(a, st);
#endif
    return cudaErrorInvalidValue;
#else
    const int cl_lat = (N + 15) / 16;
    const bool latency = lat == 1 || (lat != 0 && (long long)a.batch * cl_lat <= 148);
    if (latency) {
        if (N <= 16) return launch<Cfg<16, 1, 8>, KIND>(a, st);
        if (N <= 32) return launch<Cfg<16, 2, 8>, KIND>(a, st);
        if (N <= 64) return launch<Cfg<16, 4, 8>, KIND>(a, st);
        if (N <= 128) return launch<Cfg<16, 8, 8>, KIND>(a, st);
        if (N <= 256) return launch<Cfg<16, 16, 8>, KIND>(a, st);
        if (N <= 512) return launch<Cfg<32, 16, 8>, KIND>(a, st);
    }
    if (N <= 16) return launch<Cfg<16, 1, 8>, KIND>(a, st);
    if (N <= 32) return launch<Cfg<32, 1, 8>, KIND>(a, st);
    if (lpk == 4) {
        if (N <= 64) {
            if (KIND == K_FUSED && var2) return launch<Cfg<64, 1, 4>, K_FUSED_V2>(a, st);
            return launch<Cfg<64, 1, 4>, KIND>(a, st);
        }
        if (N <= 128) return launch<Cfg<64, 2, 4>, KIND>(a, st);
    } else {
        if (N <= 64) return launch<Cfg<64, 1, 8>, KIND>(a, st);
        if (N <= 128) return launch<Cfg<64, 2, 8>, KIND>(a, st);
    }
    if (N <= 256) return launch<Cfg<64, 4, 8>, KIND>(a, st);
    if (N <= 512) return launch<Cfg<64, 8, 8>, KIND>(a, st);
    return cudaErrorInvalidValue;
#endif
}

BatchArgs make_args(uint32_t s, uint32_t c, uint32_t N, uint32_t batch) { return make_batch_args<float>(s, c, N, batch); }

bool sizes_ok(uint32_t s, uint32_t c, uint32_t N, uint32_t batch) {
    return s == 14 && c == 7 && N >= 1 && N <= 512 && batch >= 1 && batch <= (1u << 24);
}

// reference-style comparison kernel (cooperative, one CTA per knot)
__global__ void refstyle_kernel(uint32_t s, uint32_t N, const float* d_S, const float* d_Pinv,
                                const float* d_gamma, float* d_lambda, float* d_r, float* d_p,
                                float* d_v, float* d_eta, uint32_t* d_iters, bool* d_exit,
                                uint32_t max_iter, float tol) {
    extern __shared__ float rs_smem[];
    mpcb200_refstyle::pcg_body<float>(s, N, d_S, d_Pinv, d_gamma, d_lambda, d_r, d_p, d_v, d_eta,
                                      d_iters, d_exit, max_iter, tol, rs_smem);
}

}  // namespace

extern "C" {

int mpcgpu_b200_abi_version(void) { return MPCGPU_B200_ABI_VERSION; }

const char* mpcgpu_b200_last_error(void) { return g_err.c_str(); }

uint64_t mpcgpu_b200_launch_count(void) { return g_launches.load(); }

size_t mpcgpu_b200_workspace_bytes(uint32_t s, uint32_t c, uint32_t N, uint32_t batch) {
    (void)c; (void)batch;
    // only the reference-style comparison kernel needs global scratch
    return sizeof(float) * ((size_t)2 * s * N + 2 * N);
}

int mpcgpu_b200_schur_f32(uint32_t s, uint32_t c, uint32_t N, uint32_t batch, float* d_G,
                          const float* d_C, const float* d_g, const float* d_c, float rho,
                          float* d_S, float* d_Pinv, float* d_gamma, void* stream) {
    if (!sizes_ok(s, c, N, batch)) return bad("mpcgpu_b200_schur_f32: unsupported sizes (need s=14,c=7,1<=N<=512)");
    if (!d_G || (N > 1 && !d_C) || !d_g || !d_c || !d_S || !d_Pinv || !d_gamma) return bad("mpcgpu_b200_schur_f32: null pointer");
    BatchArgs a = make_args(s, c, N, batch);
    a.G = d_G; a.C = d_C; a.g = d_g; a.c = d_c; a.rho = rho;
    a.S = d_S; a.Pinv = d_Pinv; a.gamma = d_gamma;
    cudaError_t e = dispatch<K_SCHUR>(a, (cudaStream_t)stream);
    return e == cudaSuccess ? 0 : fail(e, "mpcgpu_b200_schur_f32");
}

int mpcgpu_b200_pcg_f32(uint32_t s, uint32_t N, uint32_t batch, const float* d_S,
                        const float* d_Pinv, const float* d_gamma, float* d_lambda,
                        uint32_t max_iter, float exit_tol, uint32_t* d_iters,
                        uint8_t* d_max_iter_exit, void* d_ws, void* stream) {
    (void)d_ws;
    if (!sizes_ok(s, 7, N, batch)) return bad("mpcgpu_b200_pcg_f32: unsupported sizes (need s=14,1<=N<=512)");
    if (!d_S || !d_Pinv || !d_gamma || !d_lambda || !d_iters || !d_max_iter_exit) return bad("mpcgpu_b200_pcg_f32: null pointer");
    BatchArgs a = make_args(s, 7, N, batch);
    a.S = const_cast<float*>(d_S); a.Pinv = const_cast<float*>(d_Pinv); a.gamma = const_cast<float*>(d_gamma);
    a.lambda = d_lambda; a.max_iter = max_iter; a.exit_tol = exit_tol;
    a.iters = d_iters; a.exits = d_max_iter_exit;
    cudaError_t e = dispatch<K_PCG>(a, (cudaStream_t)stream);
    return e == cudaSuccess ? 0 : fail(e, "mpcgpu_b200_pcg_f32");
}

int mpcgpu_b200_kkt_solve_f32(uint32_t s, uint32_t c, uint32_t N, uint32_t batch, float* d_G,
                              const float* d_C, const float* d_g, const float* d_c, float rho,
                              float* d_lambda, uint32_t max_iter, float exit_tol,
                              uint32_t* d_iters, uint8_t* d_max_iter_exit, void* d_ws,
                              void* stream) {
    (void)d_ws;
    if (!sizes_ok(s, c, N, batch)) return bad("mpcgpu_b200_kkt_solve_f32: unsupported sizes (need s=14,c=7,1<=N<=512)");
    if (!d_G || (N > 1 && !d_C) || !d_g || !d_c || !d_lambda || !d_iters || !d_max_iter_exit) return bad("mpcgpu_b200_kkt_solve_f32: null pointer");
    BatchArgs a = make_args(s, c, N, batch);
    a.G = d_G; a.C = d_C; a.g = d_g; a.c = d_c; a.rho = rho;
    a.lambda = d_lambda; a.max_iter = max_iter; a.exit_tol = exit_tol;
    a.iters = d_iters; a.exits = d_max_iter_exit;
    cudaError_t e = dispatch<K_FUSED>(a, (cudaStream_t)stream);
    return e == cudaSuccess ? 0 : fail(e, "mpcgpu_b200_kkt_solve_f32");
}

int mpcgpu_b200_kkt_solve_dz_f32(uint32_t s, uint32_t c, uint32_t N, uint32_t batch, float* d_G,
                                 const float* d_C, const float* d_g, const float* d_c, float rho,
                                 float* d_lambda, float* d_dz, uint32_t max_iter, float exit_tol,
                                 uint32_t* d_iters, uint8_t* d_max_iter_exit, void* d_ws, void* stream) {
    (void)d_ws;
    if (!sizes_ok(s, c, N, batch)) return bad("mpcgpu_b200_kkt_solve_dz_f32: unsupported sizes (need s=14,c=7,1<=N<=512)");
    if (!d_G || (N > 1 && !d_C) || !d_g || !d_c || !d_lambda || !d_dz || !d_iters || !d_max_iter_exit) return bad("mpcgpu_b200_kkt_solve_dz_f32: null pointer");
    BatchArgs a = make_args(s, c, N, batch);
    a.G = d_G; a.C = d_C; a.g = d_g; a.c = d_c; a.rho = rho;
    a.lambda = d_lambda; a.max_iter = max_iter; a.exit_tol = exit_tol;
    a.iters = d_iters; a.exits = d_max_iter_exit;
    a.dz = d_dz; a.sdz = a.sg;
    cudaError_t e = dispatch<K_FUSED>(a, (cudaStream_t)stream);
    return e == cudaSuccess ? 0 : fail(e, "mpcgpu_b200_kkt_solve_dz_f32");
}

int mpcgpu_b200_kkt_solve_dz_dev_f32(uint32_t s, uint32_t c, uint32_t N, uint32_t batch, float* d_G,
                                     const float* d_C, const float* d_g, const float* d_c,
                                     const float* d_rho, float* d_lambda, float* d_dz, uint32_t max_iter,
                                     float exit_tol, uint32_t* d_iters, uint8_t* d_max_iter_exit,
                                     void* d_ws, void* stream) {
    (void)d_ws;
    if (!sizes_ok(s, c, N, batch)) return bad("mpcgpu_b200_kkt_solve_dz_dev_f32: unsupported sizes (need s=14,c=7,1<=N<=512)");
    if (!d_G || (N > 1 && !d_C) || !d_g || !d_c || !d_rho || !d_lambda || !d_dz || !d_iters || !d_max_iter_exit) return bad("mpcgpu_b200_kkt_solve_dz_dev_f32: null pointer");
    BatchArgs a = make_args(s, c, N, batch);
    a.G = d_G; a.C = d_C; a.g = d_g; a.c = d_c; a.rho = 0.f; a.rho_dev = d_rho;
    a.lambda = d_lambda; a.max_iter = max_iter; a.exit_tol = exit_tol;
    a.iters = d_iters; a.exits = d_max_iter_exit;
    a.dz = d_dz; a.sdz = a.sg;
    cudaError_t e = dispatch<K_FUSED>(a, (cudaStream_t)stream);
    return e == cudaSuccess ? 0 : fail(e, "mpcgpu_b200_kkt_solve_dz_dev_f32");
}

int mpcgpu_b200_kkt_solve_dz_rhovec_f32(uint32_t s, uint32_t c, uint32_t N, uint32_t batch, float* d_G,
                                        const float* d_C, const float* d_g, const float* d_c,
                                        const float* d_rho, float* d_lambda, float* d_dz, uint32_t max_iter,
                                        float exit_tol, uint32_t* d_iters, uint8_t* d_max_iter_exit,
                                        void* d_ws, void* stream) {
    (void)d_ws;
    if (!sizes_ok(s, c, N, batch)) return bad("mpcgpu_b200_kkt_solve_dz_rhovec_f32: unsupported sizes (need s=14,c=7,1<=N<=512)");
    if (!d_G || (N > 1 && !d_C) || !d_g || !d_c || !d_rho || !d_lambda || !d_dz || !d_iters || !d_max_iter_exit) return bad("mpcgpu_b200_kkt_solve_dz_rhovec_f32: null pointer");
    BatchArgs a = make_args(s, c, N, batch);
    a.G = d_G; a.C = d_C; a.g = d_g; a.c = d_c; a.rho = 0.f; a.rho_dev = d_rho; a.rho_stride = 1;
    a.lambda = d_lambda; a.max_iter = max_iter; a.exit_tol = exit_tol;
    a.iters = d_iters; a.exits = d_max_iter_exit;
    a.dz = d_dz; a.sdz = a.sg;
    cudaError_t e = dispatch<K_FUSED>(a, (cudaStream_t)stream);
    return e == cudaSuccess ? 0 : fail(e, "mpcgpu_b200_kkt_solve_dz_rhovec_f32");
}

int mpcgpu_b200_dz_f32(uint32_t s, uint32_t c, uint32_t N, uint32_t batch, const float* d_Ginv,
                       const float* d_C, const float* d_g, const float* d_lambda, float* d_dz, void* stream) {
    if (!sizes_ok(s, c, N, batch)) return bad("mpcgpu_b200_dz_f32: unsupported sizes (need s=14,c=7,1<=N<=512)");
    if (!d_Ginv || (N > 1 && !d_C) || !d_g || !d_lambda || !d_dz) return bad("mpcgpu_b200_dz_f32: null pointer");
    BatchArgs a = make_args(s, c, N, batch);
    a.G = const_cast<float*>(d_Ginv); a.C = d_C; a.g = d_g; a.lambda = const_cast<float*>(d_lambda);
    a.dz = d_dz; a.sdz = a.sg;
    const size_t knots = (size_t)batch * N;
    const unsigned grid = (unsigned)std::min<size_t>((knots + 31) / 32, 148u * 64u);
    dz_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(a);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(e, "mpcgpu_b200_dz_f32");
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return 0;
}

int mpcgpu_b200_pcg_refstyle_f32(uint32_t s, uint32_t N, const float* d_S, const float* d_Pinv,
                                 const float* d_gamma, float* d_lambda, uint32_t max_iter,
                                 float exit_tol, uint32_t* d_iters, uint8_t* d_max_iter_exit,
                                 void* d_ws, void* stream) {
    if (s < 1 || s > 32 || N < 1) return bad("mpcgpu_b200_pcg_refstyle_f32: unsupported sizes");
    if (!d_S || !d_Pinv || !d_gamma || !d_lambda || !d_iters || !d_max_iter_exit || !d_ws) return bad("mpcgpu_b200_pcg_refstyle_f32: null pointer");
    float* d_r = (float*)d_ws;
    float* d_p = d_r + (size_t)s * N;
    float* d_v = d_p + (size_t)s * N;
    float* d_eta = d_v + N;
    bool* d_exit = reinterpret_cast<bool*>(d_max_iter_exit);
    size_t smem = sizeof(float) * mpcb200_refstyle::smem_elems<float>(s);
    void* args[] = {&s, &N, &d_S, &d_Pinv, &d_gamma, &d_lambda, &d_r, &d_p, &d_v, &d_eta,
                    &d_iters, &d_exit, &max_iter, &exit_tol};
    cudaError_t e = cudaLaunchCooperativeKernel((void*)refstyle_kernel, dim3(N), dim3(128), args,
                                                smem, (cudaStream_t)stream);
    if (e != cudaSuccess) return fail(e, "mpcgpu_b200_pcg_refstyle_f32");
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return 0;
}

}  // extern "C"

// ---- knot-range split (configs[3]) ------------------------------------------------------------
namespace {
template <class CF>
cudaError_t launch_split(const SplitArgs& a, int ranks_here, cudaStream_t st) {
    auto kern = pcg_split_kernel<CF>;
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
    cfg.gridDim = dim3((unsigned)ranks_here * CF::CL, 1, 1);
    cfg.blockDim = dim3(CF::NT, 1, 1);
    cfg.dynamicSmemBytes = CF::SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CF::CL; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (ranks_here > 1) {   // emulated ranks wait on one another: they must be co-resident
        attr[1].id = cudaLaunchAttributeCooperative;
        attr[1].val.cooperative = 1;
        cfg.numAttrs = 2;
    }
    e = cudaLaunchKernelEx(&cfg, kern, a);
    if (e == cudaSuccess) g_launches.fetch_add(1, std::memory_order_relaxed);
    return e;
}
}  // namespace

extern "C" {

size_t mpcgpu_b200_knotsplit_mailbox_bytes(void) { return sizeof(uint2) * SPLIT_MAIL_U2; }

uint32_t mpcgpu_b200_knotsplit_epoch_stride(uint32_t max_iter) { return 2u * max_iter + 17u; }   // odd, > 2*max_iter + 8

// host-visible record of a poll that timed out: {flag, rank, expected tag, seen tag}; pinned, one per process
static unsigned* split_diag() {
    static unsigned* h = [] {
        unsigned* p = nullptr;
        if (cudaHostAlloc((void**)&p, 4 * sizeof(unsigned), cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return (unsigned*)nullptr; }
        memset(p, 0, 4 * sizeof(unsigned));
        return p;
    }();
    return h;
}

int mpcgpu_b200_knotsplit_last_timeout(uint32_t out[4]) {
    unsigned* h = split_diag();
    for (int i = 0; i < 4; i++) out[i] = h ? h[i] : 0u;
    return h && h[0] ? 1 : 0;
}

}  // extern "C"

// ---- NVLink mailbox floor: one tagged 8-byte peer store + local poll per hop, nothing else ----------------
namespace {
__global__ void pingpong_kernel(uint2* mine, uint2* peer, unsigned me, unsigned reps, unsigned tag0,
                                unsigned long long timeout_ns, unsigned* diag, unsigned long long* out_ns) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    SplitCtx X{me, 2u, tag0, nullptr, 0u, timeout_ns, diag};
    unsigned long long t0 = 0;
    for (unsigned i = 0; i <= reps; i++) {           // hop 0 is the rendezvous (the peer kernel may start later): not timed
        const unsigned tag = tag0 + i + 1u;
        if (me == 0) { sys_store_tagged(peer, (float)i, tag); (void)sys_poll_tagged(X, mine, tag); }
        else         { (void)sys_poll_tagged(X, mine, tag); sys_store_tagged(peer, (float)i, tag); }
        if (i == 0) t0 = globaltimer_ns();
    }
    *out_ns = globaltimer_ns() - t0;
}
}  // namespace

extern "C" {

// `reps` round trips between two ranks (me = 0 starts) through the LAST slot of their mailboxes; *d_out_ns = the
// kernel's own wall time over the reps (one-way latency = ns / (2 reps)).  The latency floor of one knot-split
// exchange: every PCG iteration of the split pays two such hops.
int mpcgpu_b200_knotsplit_pingpong(uint32_t me, void* d_my_mailbox, void* d_peer_mailbox, uint32_t reps, uint32_t tag0,
                                   unsigned long long* d_out_ns, void* stream) {
    if (me > 1 || !d_my_mailbox || !d_peer_mailbox || !d_out_ns || reps < 1) return bad("mpcgpu_b200_knotsplit_pingpong: bad argument");
    unsigned* diag = nullptr;
    if (unsigned* h = split_diag()) { if (cudaHostGetDevicePointer((void**)&diag, h, 0) != cudaSuccess) { cudaGetLastError(); diag = nullptr; } }
    uint2* mine = (uint2*)d_my_mailbox + SPLIT_MAIL_U2 - 1;
    uint2* peer = (uint2*)d_peer_mailbox + SPLIT_MAIL_U2 - 1;
    pingpong_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(mine, peer, me, reps, tag0, 30000000000ull, diag, d_out_ns);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(e, "mpcgpu_b200_knotsplit_pingpong");
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return 0;
}

int mpcgpu_b200_knotsplit_mailbox_create(void** d_mailbox) {
    // a dedicated cudaMalloc allocation: an IPC handle always names a whole allocation, so the
    // mailbox must not be carved out of a caching allocator's segment
    cudaError_t e = cudaMalloc(d_mailbox, sizeof(uint2) * SPLIT_MAIL_U2);
    if (e == cudaSuccess) e = cudaMemset(*d_mailbox, 0, sizeof(uint2) * SPLIT_MAIL_U2);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    return e == cudaSuccess ? 0 : fail(e, "mpcgpu_b200_knotsplit_mailbox_create");
}

int mpcgpu_b200_knotsplit_mailbox_destroy(void* d_mailbox) {
    cudaError_t e = cudaFree(d_mailbox);
    return e == cudaSuccess ? 0 : fail(e, "mpcgpu_b200_knotsplit_mailbox_destroy");
}

int mpcgpu_b200_ipc_export(const void* d_ptr, void* handle64) {
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, const_cast<void*>(d_ptr));
    if (e != cudaSuccess) return fail(e, "mpcgpu_b200_ipc_export");
    static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
    memcpy(handle64, &h, 64);
    return 0;
}

int mpcgpu_b200_ipc_open(const void* handle64, void** d_ptr) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    cudaError_t e = cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess);
    return e == cudaSuccess ? 0 : fail(e, "mpcgpu_b200_ipc_open");
}

int mpcgpu_b200_ipc_close(void* d_ptr) {
    cudaError_t e = cudaIpcCloseMemHandle(d_ptr);
    return e == cudaSuccess ? 0 : fail(e, "mpcgpu_b200_ipc_close");
}

int mpcgpu_b200_knotsplit_pcg_f32(uint32_t s, uint32_t knots_local, uint32_t ranks_here, uint32_t rank_first,
                                  uint32_t world, uint32_t epoch, void* const* d_mailboxes, const float* d_S,
                                  const float* d_Pinv, const float* d_gamma, float* d_lambda, uint32_t max_iter,
                                  float exit_tol, uint32_t* d_iters, uint8_t* d_max_iter_exit, void* stream) {
    if (s != 14 || world < 1 || world > SPLIT_MAXW || ranks_here < 1 || rank_first + ranks_here > world)
        return bad("mpcgpu_b200_knotsplit_pcg_f32: need s=14, 1 <= world <= 16, rank_first + ranks_here <= world");
    if (!d_mailboxes || !d_S || !d_Pinv || !d_gamma || !d_lambda || !d_iters || !d_max_iter_exit)
        return bad("mpcgpu_b200_knotsplit_pcg_f32: null pointer");
    SplitArgs a;
    memset(&a, 0, sizeof(a));
    a.n_local = (int)knots_local; a.grank_first = rank_first; a.world = world; a.epoch = epoch;
    // consecutive solves alternate between the two halves of every mailbox: their epochs differ by an odd stride
    a.parity = epoch & 1u;
    static const unsigned long long timeout_ms = [] {
        const char* e = getenv("MPCGPU_B200_KNOTSPLIT_TIMEOUT_MS");
        const long long v = e ? atoll(e) : 0;
        return (unsigned long long)(v > 0 ? v : 30000);
    }();
    a.timeout_ns = timeout_ms * 1000000ull;
    a.diag = nullptr;
    if (unsigned* h = split_diag()) {
        unsigned* d = nullptr;
        if (cudaHostGetDevicePointer((void**)&d, h, 0) == cudaSuccess) a.diag = d; else cudaGetLastError();
    }
    for (uint32_t r = 0; r < world; r++) {
        if (!d_mailboxes[r]) return bad("mpcgpu_b200_knotsplit_pcg_f32: null mailbox");
        a.mail[r] = (uint2*)d_mailboxes[r];
    }
    a.S = d_S; a.Pinv = d_Pinv; a.gamma = d_gamma; a.lambda = d_lambda;
    a.max_iter = max_iter; a.exit_tol = exit_tol; a.iters = d_iters; a.exits = d_max_iter_exit;
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t e;
    switch (knots_local) {
#ifndef MPCB200_MINIMAL
        case 16: e = launch_split<Cfg<16, 1, 8>>(a, (int)ranks_here, st); break;
        case 32: e = launch_split<Cfg<16, 2, 8>>(a, (int)ranks_here, st); break;
        case 64: e = launch_split<Cfg<16, 4, 8>>(a, (int)ranks_here, st); break;
        case 128: e = launch_split<Cfg<16, 8, 8>>(a, (int)ranks_here, st); break;
        case 256: e = launch_split<Cfg<16, 16, 8>>(a, (int)ranks_here, st); break;
        case 512: e = launch_split<Cfg<32, 16, 8>>(a, (int)ranks_here, st); break;
#endif
        default: return bad("mpcgpu_b200_knotsplit_pcg_f32: knots per rank must be 16, 32, 64, 128, 256 or 512");
    }
    return e == cudaSuccess ? 0 : fail(e, "mpcgpu_b200_knotsplit_pcg_f32");
}

}  // extern "C"
