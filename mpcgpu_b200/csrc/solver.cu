// solver.cu -- host-buffer solver object of the C ABI (include/mpcgpu_b200.h).
//
// Persistent device buffers + two streams; a batch is cut into chunks and the H2D copy of
// chunk i+1 overlaps the fused KKT kernel of chunk i and the D2H copy of chunk i-1.
// Nothing is allocated per solve (the reference slot allocates 19 buffers, 8 streams and a
// cuBLAS handle per sqpSolve call: include/pcg/sqp.cuh:75-148,429-457).
#include "../../include/mpcgpu_b200.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cstring>
#include <new>
#include <string>

#include "capi_internal.h"

struct mpcgpu_b200_solver {
    int device;
    uint32_t s, c, N, max_batch;
    size_t nG, nC, ng, nc, nv;   // elements per instance
    uint32_t chunk;              // instances per pipeline stage
    static constexpr int NBUF = 3;
    float *dG[NBUF], *dC[NBUF], *dg[NBUF], *dc[NBUF], *dlam[NBUF];
    uint32_t* diters[NBUF];
    uint8_t* dexit[NBUF];
    cudaStream_t st_in, st_k, st_out;
    cudaEvent_t ev_in[NBUF], ev_k[NBUF], ev_out[NBUF];
};

namespace {
// every failure leaves its text in mpcgpu_b200_last_error(), like the other entry points of the ABI
int sfail(cudaError_t e, const char* what = "mpcgpu_b200_solver") { return mpcb200::detail::fail(e, what); }
int sbad(const char* what) { return mpcb200::detail::bad(what); }
}  // namespace

extern "C" {

int mpcgpu_b200_solver_create(mpcgpu_b200_solver** out, int device, uint32_t s, uint32_t c,
                              uint32_t N, uint32_t max_batch) {
    if (!out || s != 14 || c != 7 || N < 1 || N > 512 || max_batch < 1)
        return sbad("mpcgpu_b200_solver_create: need out != NULL, s=14, c=7, 1<=N<=512, max_batch>=1");
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return sfail(e, "mpcgpu_b200_solver_create: cudaSetDevice");
    mpcgpu_b200_solver* S = new (std::nothrow) mpcgpu_b200_solver();
    if (!S) return sfail(cudaErrorMemoryAllocation, "mpcgpu_b200_solver_create");
    memset(S, 0, sizeof(*S));
    S->device = device; S->s = s; S->c = c; S->N = N; S->max_batch = max_batch;
    S->nG = (size_t)(s * s + c * c) * N - c * c;
    S->nC = (size_t)(s * s + s * c) * (N - 1);
    S->ng = (size_t)(s + c) * N - c;
    S->nc = (size_t)s * N;
    S->nv = (size_t)s * N;
    // chunk: enough clusters to fill the GPU a few times over, small enough to pipeline
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    uint32_t chunk = (uint32_t)std::max(1, sms * 4);
    S->chunk = std::min(chunk, max_batch);
    for (int b = 0; b < mpcgpu_b200_solver::NBUF && e == cudaSuccess; b++) {
        const size_t k = S->chunk;
        if (e == cudaSuccess) e = cudaMalloc(&S->dG[b], sizeof(float) * S->nG * k);
        if (e == cudaSuccess) e = cudaMalloc(&S->dC[b], sizeof(float) * std::max<size_t>(S->nC, 1) * k);
        if (e == cudaSuccess) e = cudaMalloc(&S->dg[b], sizeof(float) * S->ng * k);
        if (e == cudaSuccess) e = cudaMalloc(&S->dc[b], sizeof(float) * S->nc * k);
        if (e == cudaSuccess) e = cudaMalloc(&S->dlam[b], sizeof(float) * S->nv * k);
        if (e == cudaSuccess) e = cudaMalloc(&S->diters[b], sizeof(uint32_t) * k);
        if (e == cudaSuccess) e = cudaMalloc(&S->dexit[b], k);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&S->ev_in[b], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&S->ev_k[b], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&S->ev_out[b], cudaEventDisableTiming);
    }
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&S->st_in, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&S->st_k, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&S->st_out, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        mpcgpu_b200_solver_destroy(S);
        return sfail(e, "mpcgpu_b200_solver_create: device buffers / streams");
    }
    *out = S;
    return 0;
}

void mpcgpu_b200_solver_destroy(mpcgpu_b200_solver* S) {
    if (!S) return;
    cudaSetDevice(S->device);
    for (int b = 0; b < mpcgpu_b200_solver::NBUF; b++) {
        cudaFree(S->dG[b]); cudaFree(S->dC[b]); cudaFree(S->dg[b]); cudaFree(S->dc[b]);
        cudaFree(S->dlam[b]); cudaFree(S->diters[b]); cudaFree(S->dexit[b]);
        if (S->ev_in[b]) cudaEventDestroy(S->ev_in[b]);
        if (S->ev_k[b]) cudaEventDestroy(S->ev_k[b]);
        if (S->ev_out[b]) cudaEventDestroy(S->ev_out[b]);
    }
    if (S->st_in) cudaStreamDestroy(S->st_in);
    if (S->st_k) cudaStreamDestroy(S->st_k);
    if (S->st_out) cudaStreamDestroy(S->st_out);
    delete S;
}

int mpcgpu_b200_solver_solve_host_f32(mpcgpu_b200_solver* S, uint32_t batch, const float* h_G,
                                      const float* h_C, const float* h_g, const float* h_c,
                                      float rho, float* h_lambda, uint32_t max_iter,
                                      float exit_tol, uint32_t* h_iters, uint8_t* h_exit,
                                      float* h_Ginv) {
    return mpcgpu_b200_solver_solve_host_io_f32(S, batch, h_G, h_C, h_g, h_c, rho, h_lambda, h_lambda, max_iter,
                                                exit_tol, h_iters, h_exit, h_Ginv);
}

int mpcgpu_b200_solver_solve_host_io_f32(mpcgpu_b200_solver* S, uint32_t batch, const float* h_G,
                                         const float* h_C, const float* h_g, const float* h_c,
                                         float rho, const float* h_lambda0, float* h_lambda,
                                         uint32_t max_iter, float exit_tol, uint32_t* h_iters,
                                         uint8_t* h_exit, float* h_Ginv) {
    if (!S || batch < 1 || batch > S->max_batch || !h_G || (S->N > 1 && !h_C) || !h_g || !h_c ||
        !h_lambda || !h_iters || !h_exit)
        return sbad("mpcgpu_b200_solver_solve_host: null pointer or batch outside [1, max_batch]");
    cudaError_t e = cudaSetDevice(S->device);
    if (e != cudaSuccess) return sfail(e, "mpcgpu_b200_solver_solve_host: cudaSetDevice");
    constexpr int NBUF = mpcgpu_b200_solver::NBUF;
    const uint32_t nchunks = (batch + S->chunk - 1) / S->chunk;
#define CK(x) do { e = (x); if (e != cudaSuccess) return sfail(e, "mpcgpu_b200_solver_solve_host: " #x); } while (0)
    for (uint32_t ci = 0; ci < nchunks; ci++) {
        const int b = ci % NBUF;
        const size_t first = (size_t)ci * S->chunk;
        const size_t k = std::min<size_t>(S->chunk, batch - first);
        // buffer b is free once its previous D2H has finished
        if (ci >= (uint32_t)NBUF) CK(cudaStreamWaitEvent(S->st_in, S->ev_out[b], 0));
        CK(cudaMemcpyAsync(S->dG[b], h_G + first * S->nG, sizeof(float) * S->nG * k, cudaMemcpyHostToDevice, S->st_in));
        if (S->nC) CK(cudaMemcpyAsync(S->dC[b], h_C + first * S->nC, sizeof(float) * S->nC * k, cudaMemcpyHostToDevice, S->st_in));
        CK(cudaMemcpyAsync(S->dg[b], h_g + first * S->ng, sizeof(float) * S->ng * k, cudaMemcpyHostToDevice, S->st_in));
        CK(cudaMemcpyAsync(S->dc[b], h_c + first * S->nc, sizeof(float) * S->nc * k, cudaMemcpyHostToDevice, S->st_in));
        if (h_lambda0) CK(cudaMemcpyAsync(S->dlam[b], h_lambda0 + first * S->nv, sizeof(float) * S->nv * k, cudaMemcpyHostToDevice, S->st_in));
        else CK(cudaMemsetAsync(S->dlam[b], 0, sizeof(float) * S->nv * k, S->st_in));   // cold start
        CK(cudaEventRecord(S->ev_in[b], S->st_in));
        CK(cudaStreamWaitEvent(S->st_k, S->ev_in[b], 0));
        int rc = mpcgpu_b200_kkt_solve_f32(S->s, S->c, S->N, (uint32_t)k, S->dG[b], S->dC[b], S->dg[b],
                                           S->dc[b], rho, S->dlam[b], max_iter, exit_tol,
                                           S->diters[b], S->dexit[b], nullptr, (void*)S->st_k);
        if (rc) return rc;
        CK(cudaEventRecord(S->ev_k[b], S->st_k));
        CK(cudaStreamWaitEvent(S->st_out, S->ev_k[b], 0));
        CK(cudaMemcpyAsync(h_lambda + first * S->nv, S->dlam[b], sizeof(float) * S->nv * k, cudaMemcpyDeviceToHost, S->st_out));
        CK(cudaMemcpyAsync(h_iters + first, S->diters[b], sizeof(uint32_t) * k, cudaMemcpyDeviceToHost, S->st_out));
        CK(cudaMemcpyAsync(h_exit + first, S->dexit[b], k, cudaMemcpyDeviceToHost, S->st_out));
        if (h_Ginv) CK(cudaMemcpyAsync(h_Ginv + first * S->nG, S->dG[b], sizeof(float) * S->nG * k, cudaMemcpyDeviceToHost, S->st_out));
        CK(cudaEventRecord(S->ev_out[b], S->st_out));
    }
    CK(cudaStreamSynchronize(S->st_out));
#undef CK
    return 0;
}

}  // extern "C"
