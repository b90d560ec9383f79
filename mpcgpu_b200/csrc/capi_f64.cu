// capi_f64.cu -- fp64 twins of the C ABI (include/mpcgpu_b200.h), for builds of the reference with
// USE_DOUBLES 1 (linsys_t = double, include/common/settings.cuh:49-57).
// The kernels are the templates of kkt_device.cuh / kernels.cuh instantiated for double:
//   Cfg<16, CL, 8, 0, double>: 16 knots per 128-thread CTA, two rows per lane (8 lanes per knot), S rows
//   (168 registers) stay in registers, Pinv rows in shared memory (no tensor memory: its columns are 32 bit
//   and the rows would not fit beside a second copy), pair-wise DFMA, the same st.async + mbarrier cluster
//   reductions carrying 8-byte partials.  N <= 256 (cluster of 16).  Batches of more than one wave use
//   Cfg<8, CL, 8, 0, double> (three 64-thread CTAs per SM); MPCGPU_B200_F64_THROUGHPUT_TILING=0 switches it off.
#include "../../include/mpcgpu_b200.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "launch.cuh"

using namespace mpcb200;
using detail::bad;
using detail::fail;

namespace {

using Args = BatchArgsT<double>;

template <int CL> using D8 = Cfg<16, CL, 8, 0, double>;     // latency tiling: 16 knots per 128-thread CTA, two rows per lane
template <int CL> using T8 = Cfg<8, CL, 8, 0, double>;      // throughput tiling: 8 knots per 64-thread CTA

// Measured on B200 (tools/f64_ab.py, fused solves/s, 1024 x N=64): 16 knots x 16 lanes 148 k, 16 knots x 8 lanes
// 162 k, 12 knots x 8 lanes (two CTAs per SM) 187 k, 8 knots x 8 lanes (69 KB of shared memory: THREE CTAs per SM
// hide each other's reduction latency) 208 k; the same order at N = 32 (1.15 M) and N = 128 (108 k).
template <int KIND>
cudaError_t dispatch(const Args& a, cudaStream_t st) {
    const int N = a.N;
    static const int thr = [] { const char* e = getenv("MPCGPU_B200_F64_THROUGHPUT_TILING"); return e ? atoi(e) : 1; }();
    if (thr) {      // batches of more than one wave
        if (N > 16 && N <= 32 && (long long)a.batch * 4 > 148) return launch<T8<4>, KIND>(a, st);
        if (N > 32 && N <= 64 && (long long)a.batch * 8 > 148) return launch<T8<8>, KIND>(a, st);
        if (N > 64 && N <= 128 && (long long)a.batch * 16 > 148) return launch<T8<16>, KIND>(a, st);
    }
    if (N <= 16) return launch<D8<1>, KIND>(a, st);
    if (N <= 32) return launch<D8<2>, KIND>(a, st);
    if (N <= 64) return launch<D8<4>, KIND>(a, st);
    if (N <= 128) return launch<D8<8>, KIND>(a, st);
    if (N <= 256) return launch<D8<16>, KIND>(a, st);
    return cudaErrorInvalidValue;
}

Args make_args(uint32_t s, uint32_t c, uint32_t N, uint32_t batch) { return make_batch_args<double>(s, c, N, batch); }

bool sizes_ok(uint32_t s, uint32_t c, uint32_t N, uint32_t batch) {
    return s == 14 && c == 7 && N >= 1 && N <= 256 && batch >= 1 && batch <= (1u << 24);
}

}  // namespace

extern "C" {

int mpcgpu_b200_schur_f64(uint32_t s, uint32_t c, uint32_t N, uint32_t batch, double* d_G,
                          const double* d_C, const double* d_g, const double* d_c, double rho,
                          double* d_S, double* d_Pinv, double* d_gamma, void* stream) {
    if (!sizes_ok(s, c, N, batch)) return bad("mpcgpu_b200_schur_f64: unsupported sizes (need s=14,c=7,1<=N<=256)");
    if (!d_G || (N > 1 && !d_C) || !d_g || !d_c || !d_S || !d_Pinv || !d_gamma) return bad("mpcgpu_b200_schur_f64: null pointer");
    Args a = make_args(s, c, N, batch);
    a.G = d_G; a.C = d_C; a.g = d_g; a.c = d_c; a.rho = rho;
    a.S = d_S; a.Pinv = d_Pinv; a.gamma = d_gamma;
    cudaError_t e = dispatch<K_SCHUR>(a, (cudaStream_t)stream);
    return e == cudaSuccess ? 0 : fail(e, "mpcgpu_b200_schur_f64");
}

int mpcgpu_b200_pcg_f64(uint32_t s, uint32_t N, uint32_t batch, const double* d_S, const double* d_Pinv,
                        const double* d_gamma, double* d_lambda, uint32_t max_iter, double exit_tol,
                        uint32_t* d_iters, uint8_t* d_max_iter_exit, void* d_ws, void* stream) {
    (void)d_ws;
    if (!sizes_ok(s, 7, N, batch)) return bad("mpcgpu_b200_pcg_f64: unsupported sizes (need s=14,1<=N<=256)");
    if (!d_S || !d_Pinv || !d_gamma || !d_lambda || !d_iters || !d_max_iter_exit) return bad("mpcgpu_b200_pcg_f64: null pointer");
    Args a = make_args(s, 7, N, batch);
    a.S = const_cast<double*>(d_S); a.Pinv = const_cast<double*>(d_Pinv); a.gamma = const_cast<double*>(d_gamma);
    a.lambda = d_lambda; a.max_iter = max_iter; a.exit_tol = exit_tol;
    a.iters = d_iters; a.exits = d_max_iter_exit;
    cudaError_t e = dispatch<K_PCG>(a, (cudaStream_t)stream);
    return e == cudaSuccess ? 0 : fail(e, "mpcgpu_b200_pcg_f64");
}

int mpcgpu_b200_kkt_solve_dz_f64(uint32_t s, uint32_t c, uint32_t N, uint32_t batch, double* d_G,
                                 const double* d_C, const double* d_g, const double* d_c, double rho,
                                 double* d_lambda, double* d_dz, uint32_t max_iter, double exit_tol,
                                 uint32_t* d_iters, uint8_t* d_max_iter_exit, void* d_ws, void* stream) {
    (void)d_ws;
    if (!sizes_ok(s, c, N, batch)) return bad("mpcgpu_b200_kkt_solve_f64: unsupported sizes (need s=14,c=7,1<=N<=256)");
    if (!d_G || (N > 1 && !d_C) || !d_g || !d_c || !d_lambda || !d_iters || !d_max_iter_exit) return bad("mpcgpu_b200_kkt_solve_f64: null pointer");
    Args a = make_args(s, c, N, batch);
    a.G = d_G; a.C = d_C; a.g = d_g; a.c = d_c; a.rho = rho;
    a.lambda = d_lambda; a.max_iter = max_iter; a.exit_tol = exit_tol;
    a.iters = d_iters; a.exits = d_max_iter_exit;
    a.dz = d_dz; a.sdz = a.sg;
    cudaError_t e = dispatch<K_FUSED>(a, (cudaStream_t)stream);
    return e == cudaSuccess ? 0 : fail(e, "mpcgpu_b200_kkt_solve_f64");
}

int mpcgpu_b200_kkt_solve_dz_dev_f64(uint32_t s, uint32_t c, uint32_t N, uint32_t batch, double* d_G,
                                     const double* d_C, const double* d_g, const double* d_c, const double* d_rho,
                                     double* d_lambda, double* d_dz, uint32_t max_iter, double exit_tol,
                                     uint32_t* d_iters, uint8_t* d_max_iter_exit, void* d_ws, void* stream) {
    (void)d_ws;
    if (!sizes_ok(s, c, N, batch)) return bad("mpcgpu_b200_kkt_solve_dz_dev_f64: unsupported sizes (need s=14,c=7,1<=N<=256)");
    if (!d_G || (N > 1 && !d_C) || !d_g || !d_c || !d_rho || !d_lambda || !d_dz || !d_iters || !d_max_iter_exit) return bad("mpcgpu_b200_kkt_solve_dz_dev_f64: null pointer");
    Args a = make_args(s, c, N, batch);
    a.G = d_G; a.C = d_C; a.g = d_g; a.c = d_c; a.rho = 0.0; a.rho_dev = d_rho;
    a.lambda = d_lambda; a.max_iter = max_iter; a.exit_tol = exit_tol;
    a.iters = d_iters; a.exits = d_max_iter_exit;
    a.dz = d_dz; a.sdz = a.sg;
    cudaError_t e = dispatch<K_FUSED>(a, (cudaStream_t)stream);
    return e == cudaSuccess ? 0 : fail(e, "mpcgpu_b200_kkt_solve_dz_dev_f64");
}

int mpcgpu_b200_kkt_solve_f64(uint32_t s, uint32_t c, uint32_t N, uint32_t batch, double* d_G,
                              const double* d_C, const double* d_g, const double* d_c, double rho,
                              double* d_lambda, uint32_t max_iter, double exit_tol, uint32_t* d_iters,
                              uint8_t* d_max_iter_exit, void* d_ws, void* stream) {
    return mpcgpu_b200_kkt_solve_dz_f64(s, c, N, batch, d_G, d_C, d_g, d_c, rho, d_lambda, nullptr, max_iter,
                                        exit_tol, d_iters, d_max_iter_exit, d_ws, stream);
}

int mpcgpu_b200_dz_f64(uint32_t s, uint32_t c, uint32_t N, uint32_t batch, const double* d_Ginv,
                       const double* d_C, const double* d_g, const double* d_lambda, double* d_dz, void* stream) {
    if (!sizes_ok(s, c, N, batch)) return bad("mpcgpu_b200_dz_f64: unsupported sizes (need s=14,c=7,1<=N<=256)");
    if (!d_Ginv || (N > 1 && !d_C) || !d_g || !d_lambda || !d_dz) return bad("mpcgpu_b200_dz_f64: null pointer");
    Args a = make_args(s, c, N, batch);
    a.G = const_cast<double*>(d_Ginv); a.C = d_C; a.g = d_g; a.lambda = const_cast<double*>(d_lambda);
    a.dz = d_dz; a.sdz = a.sg;
    const size_t knots = (size_t)batch * N;
    const unsigned grid = (unsigned)std::min<size_t>((knots + 31) / 32, 148u * 64u);
    dz_kernel<double><<<grid, 256, 0, (cudaStream_t)stream>>>(a);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(e, "mpcgpu_b200_dz_f64");
    detail::count_launch();
    return 0;
}

}  // extern "C"
