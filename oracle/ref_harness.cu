// ref_harness.cu -- TEST INFRASTRUCTURE ONLY.
//
// Command-line harness around the REFERENCE's own device code, compiled from the headers where
// they lie under /root/reference (never copied) against the drop-in shim headers of
// mpcgpu_b200/shim (which stand in for the two submodules that are empty in the checkout).
// Built by oracle/build.py into oracle/_ref/ref_harness so that it travels to the GPU box.
//
//   ref_harness kkt   <N> <row0> <perturb> <out.bin>
//        the reference's generate_kkt_submatrices (include/common/kkt.cuh:22-163, launched as
//        include/pcg/sqp.cuh:212) on rows [row0, row0+N) of examples/trajfiles/0_0_* ->
//        out.bin = [G | C | g | c] float32: REAL IIWA KKT blocks (GRiD dynamics).
//   ref_harness schur <N> <rho> <in.bin> <out.bin>
//        the reference's OWN form_schur_system<float> (include/pcg/linsys_setup.cuh:620-657)
//        on in.bin = [G | C | g | c] -> out.bin = [S | Pinv | gamma | Ginv]
//   ref_harness schur_shim <N> <rho> <in.bin> <out.bin>
//        the same through the header shim's stage 1 (mpcb200_shim::form_schur_system_n, the function the
//        unchanged include/pcg/sqp.cuh:243 binds to under the L1 shim): same file formats
//   ref_harness pcg   <N> <tol> <max_iter> <in.bin> <out.bin>
//        the header shim's pcg<float,14,N>, launched EXACTLY like include/pcg/sqp.cuh:275
//        (cooperative, grid = N, block = PCG_NUM_THREADS, pcgSharedMemSize) on
//        in.bin = [S | Pinv | gamma | lambda0] -> out.bin = [lambda | iters | max_iter_exit]
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
#define MPCGPU_B200_SHIM_KEEP_REFERENCE_SCHUR 1   // form_schur_system<float> below must stay the REFERENCE's kernel
#include "gpu_pcg.cuh"
#include "common/kkt.cuh"            // (also defines the cgrps alias linsys_setup.cuh relies on)
#include "pcg/linsys_setup.cuh"
#include "utils/experiment.cuh"

static std::vector<float> read_bin(const char* path) {
    FILE* f = fopen(path, "rb");
    if (!f) { fprintf(stderr, "cannot open %s\n", path); exit(2); }
    fseek(f, 0, SEEK_END);
    long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    std::vector<float> v(n / 4);
    if (fread(v.data(), 4, v.size(), f) != v.size()) { fprintf(stderr, "short read %s\n", path); exit(2); }
    fclose(f);
    return v;
}
static void write_bin(const char* path, const std::vector<float>& v) {
    FILE* f = fopen(path, "wb");
    if (!f) { fprintf(stderr, "cannot write %s\n", path); exit(2); }
    fwrite(v.data(), 4, v.size(), f);
    fclose(f);
}
template <typename T>
static T* to_dev(const T* h, size_t n) {
    T* d;
    gpuErrchk(cudaMalloc(&d, std::max<size_t>(n, 1) * sizeof(T)));
    if (n) gpuErrchk(cudaMemcpy(d, h, n * sizeof(T), cudaMemcpyHostToDevice));
    return d;
}
static void from_dev(std::vector<float>& out, const float* d, size_t n) {
    const size_t o = out.size();
    out.resize(o + n);
    gpuErrchk(cudaMemcpy(out.data() + o, d, n * sizeof(float), cudaMemcpyDeviceToHost));
}

struct Sizes {
    size_t G, C, g, c, S, v;
    explicit Sizes(uint32_t N) {
        const uint32_t s = 14, m = 7;
        G = (size_t)(s * s + m * m) * N - m * m; C = (size_t)(s * s + s * m) * (N - 1);
        g = (size_t)(s + m) * N - m; c = (size_t)s * N; S = (size_t)3 * s * s * N; v = (size_t)s * N;
    }
};

static int do_kkt(uint32_t N, uint32_t row0, float perturb, const char* out) {
    const uint32_t s = 14, m = 7;
    auto ee = readCSVToVecVec<float>("examples/trajfiles/0_0_eepos.traj");
    auto xu = readCSVToVecVec<float>("examples/trajfiles/0_0_traj.csv");
    if (ee.size() < row0 + N || xu.size() < row0 + N) { fprintf(stderr, "trajectory too short\n"); return 3; }
    std::vector<float> h_ee, h_xu;
    for (uint32_t k = 0; k < N; k++) {
        h_ee.insert(h_ee.end(), ee[row0 + k].begin(), ee[row0 + k].begin() + 6);
        const size_t take = (k + 1 < N) ? s + m : s;
        h_xu.insert(h_xu.end(), xu[row0 + k].begin(), xu[row0 + k].begin() + take);
    }
    std::vector<float> h_xs(h_xu.begin(), h_xu.begin() + s);
    // deterministic perturbation of the iterate (an SQP iterate is not the nominal trajectory:
    // non-zero integrator defects and cost gradients)
    for (size_t i = 0; i < h_xu.size(); i++) h_xu[i] += perturb * std::sin(0.37f * (float)i + 1.3f);
    Sizes z(N);
    float *d_G, *d_C, *d_g, *d_c;
    gpuErrchk(cudaMalloc(&d_G, z.G * 4)); gpuErrchk(cudaMalloc(&d_C, z.C * 4));
    gpuErrchk(cudaMalloc(&d_g, z.g * 4)); gpuErrchk(cudaMalloc(&d_c, z.c * 4));
    gpuErrchk(cudaMemset(d_c, 0, z.c * 4));
    float* d_ee = to_dev(h_ee.data(), h_ee.size());
    float* d_xu = to_dev(h_xu.data(), h_xu.size());
    float* d_xs = to_dev(h_xs.data(), h_xs.size());
    void* d_dyn = gato_plant::initializeDynamicsConstMem<float>();
    const float timestep = 0.015625f;                                   // examples/track_iiwa_pcg.cu:19
    generate_kkt_submatrices<float><<<N, KKT_THREADS, 2 * get_kkt_smem_size<float>(s, m)>>>(
        s, m, N, d_G, d_C, d_g, d_c, d_dyn, timestep, d_ee, d_xs, d_xu);  // include/pcg/sqp.cuh:212
    gpuErrchk(cudaPeekAtLastError());
    gpuErrchk(cudaDeviceSynchronize());
    std::vector<float> o;
    from_dev(o, d_G, z.G); from_dev(o, d_C, z.C); from_dev(o, d_g, z.g); from_dev(o, d_c, z.c);
    write_bin(out, o);
    return 0;
}

template <uint32_t N>
static void shim_schur_n(float* d_G, float* d_C, float* d_g, float* d_c, float* d_S, float* d_P, float* d_gamma, float rho) {
    mpcb200_shim::form_schur_system_n<float, N>(d_G, d_C, d_g, d_c, d_S, d_P, d_gamma, rho);
}

static int do_schur(uint32_t N, float rho, const char* in, const char* out, bool shim = false) {
    Sizes z(N);
    auto h = read_bin(in);
    if (h.size() != z.G + z.C + z.g + z.c) { fprintf(stderr, "bad input size\n"); return 3; }
    float* d_G = to_dev(h.data(), z.G);
    float* d_C = to_dev(h.data() + z.G, z.C);
    float* d_g = to_dev(h.data() + z.G + z.C, z.g);
    float* d_c = to_dev(h.data() + z.G + z.C + z.g, z.c);
    float *d_S, *d_P, *d_gamma;
    gpuErrchk(cudaMalloc(&d_S, z.S * 4)); gpuErrchk(cudaMalloc(&d_P, z.S * 4)); gpuErrchk(cudaMalloc(&d_gamma, z.v * 4));
    gpuErrchk(cudaMemset(d_S, 0, z.S * 4)); gpuErrchk(cudaMemset(d_P, 0, z.S * 4));   // corner blocks are never written
    if (!shim) form_schur_system<float>(14, 7, N, d_G, d_C, d_g, d_c, d_S, d_P, d_gamma, rho);   // the reference's own kernel
    else switch (N) {
        case 8: shim_schur_n<8>(d_G, d_C, d_g, d_c, d_S, d_P, d_gamma, rho); break;
        case 32: shim_schur_n<32>(d_G, d_C, d_g, d_c, d_S, d_P, d_gamma, rho); break;
        case 64: shim_schur_n<64>(d_G, d_C, d_g, d_c, d_S, d_P, d_gamma, rho); break;
        case 128: shim_schur_n<128>(d_G, d_C, d_g, d_c, d_S, d_P, d_gamma, rho); break;
        case 512: shim_schur_n<512>(d_G, d_C, d_g, d_c, d_S, d_P, d_gamma, rho); break;
        default: fprintf(stderr, "schur_shim: N must be 8, 32, 64, 128 or 512\n"); return 3;
    }
    gpuErrchk(cudaPeekAtLastError());
    gpuErrchk(cudaDeviceSynchronize());
    std::vector<float> o;
    from_dev(o, d_S, z.S); from_dev(o, d_P, z.S); from_dev(o, d_gamma, z.v); from_dev(o, d_G, z.G);
    write_bin(out, o);
    return 0;
}

template <uint32_t N>
static int do_pcg_n(float tol, uint32_t max_iter, const char* in, const char* out) {
    Sizes z(N);
    auto h = read_bin(in);
    if (h.size() != 2 * z.S + 2 * z.v) { fprintf(stderr, "bad input size\n"); return 3; }
    float* d_S = to_dev(h.data(), z.S);
    float* d_Pinv = to_dev(h.data() + z.S, z.S);
    float* d_gamma = to_dev(h.data() + 2 * z.S, z.v);
    float* d_lambda = to_dev(h.data() + 2 * z.S + z.v, z.v);
    float *d_r, *d_p, *d_v_temp, *d_eta_new_temp;
    gpuErrchk(cudaMalloc(&d_r, z.v * 4)); gpuErrchk(cudaMalloc(&d_p, z.v * 4));
    gpuErrchk(cudaMalloc(&d_v_temp, N * 4)); gpuErrchk(cudaMalloc(&d_eta_new_temp, N * 4));
    uint32_t* d_pcg_iters; bool* d_pcg_exit;
    gpuErrchk(cudaMalloc(&d_pcg_iters, sizeof(uint32_t))); gpuErrchk(cudaMalloc(&d_pcg_exit, sizeof(bool)));
    pcg_config<float> config;
    config.pcg_exit_tol = tol; config.pcg_max_iter = max_iter;
    uint32_t knot_points = N, state_size = 14;
    checkPcgOccupancy<float>((void*)pcg<float, 14, N>, PCG_NUM_THREADS, state_size, knot_points);   // examples/track_iiwa_pcg.cu:24
    // --- verbatim shape of include/pcg/sqp.cuh:142-166,275-277 ---
    void* pcg_kernel = (void*)pcg<float, 14, N>;
    void* pcgKernelArgs[] = {(void*)&d_S, (void*)&d_Pinv, (void*)&d_gamma, (void*)&d_lambda, (void*)&d_r, (void*)&d_p,
                             (void*)&d_v_temp, (void*)&d_eta_new_temp, (void*)&d_pcg_iters, (void*)&d_pcg_exit,
                             (void*)&config.pcg_max_iter, (void*)&config.pcg_exit_tol};
    size_t ppcg_kernel_smem_size = pcgSharedMemSize<float>(state_size, knot_points);
    gpuErrchk(cudaLaunchCooperativeKernel(pcg_kernel, knot_points, PCG_NUM_THREADS, pcgKernelArgs, ppcg_kernel_smem_size));
    uint32_t pcg_iters; bool pcg_exit;
    gpuErrchk(cudaMemcpy(&pcg_iters, d_pcg_iters, sizeof(uint32_t), cudaMemcpyDeviceToHost));
    gpuErrchk(cudaMemcpy(&pcg_exit, d_pcg_exit, sizeof(bool), cudaMemcpyDeviceToHost));
    gpuErrchk(cudaPeekAtLastError());
    std::vector<float> o;
    from_dev(o, d_lambda, z.v);
    o.push_back((float)pcg_iters); o.push_back(pcg_exit ? 1.f : 0.f);
    write_bin(out, o);
    return 0;
}

int main(int argc, char** argv) {
    if (argc >= 6 && !strcmp(argv[1], "kkt")) return do_kkt(atoi(argv[2]), atoi(argv[3]), atof(argv[4]), argv[5]);
    if (argc >= 6 && !strcmp(argv[1], "schur")) return do_schur(atoi(argv[2]), atof(argv[3]), argv[4], argv[5]);
    if (argc >= 6 && !strcmp(argv[1], "schur_shim")) return do_schur(atoi(argv[2]), atof(argv[3]), argv[4], argv[5], true);
    if (argc >= 7 && !strcmp(argv[1], "pcg")) {
        const int N = atoi(argv[2]); const float tol = atof(argv[3]); const uint32_t mi = atoi(argv[4]);
        switch (N) {
            case 8: return do_pcg_n<8>(tol, mi, argv[5], argv[6]);
            case 32: return do_pcg_n<32>(tol, mi, argv[5], argv[6]);
            case 64: return do_pcg_n<64>(tol, mi, argv[5], argv[6]);
            case 128: return do_pcg_n<128>(tol, mi, argv[5], argv[6]);
            case 512: return do_pcg_n<512>(tol, mi, argv[5], argv[6]);
            default: fprintf(stderr, "pcg: N must be 8, 32, 64, 128 or 512\n"); return 3;
        }
    }
    fprintf(stderr, "usage: ref_harness kkt|schur|pcg ... (see the header of oracle/ref_harness.cu)\n");
    return 1;
}
