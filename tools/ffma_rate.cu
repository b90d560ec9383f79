// ffma_rate.cu -- what the FP32 pipe of one SM really issues per clock: scalar FFMA against FFMA2 (fma.rn.f32x2, two
// FMAs per instruction), 16 independent accumulator chains per thread, 4 / 8 / 16 / 32 warps per SM, in two operand
// patterns: acc = acc * b + c (operand-reuse friendly) and the matvec-like acc_i = coef_i * x + acc_i (16 distinct
// coefficient registers per thread, like the S / Pinv rows of the fused KKT kernel).  The fused kernel's matvecs are
// FFMA2 streams whose stall reason is `math pipe throttle` at ~4.4 cycles per FFMA2 per scheduler
// (profiles/r2b_fused_loop_by_instruction.txt): this measures the ceiling that goes with it.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 tools/ffma_rate.cu -o tools/ffma_rate.bin
#include <cstdio>
#include <cuda_runtime.h>

constexpr int CH = 16, ITERS = 4096;

template <int MODE>   // 0: FFMA  1: FFMA2 (acc = acc * b + c);  2: FFMA  3: FFMA2 (acc_i = coef_i * x + acc_i)
__global__ void rate(float* out, long long* cycles, const float* __restrict__ coef) {
    const float b = coef[threadIdx.x & 1], c = coef[2 + (threadIdx.x & 1)];   // from memory: register operands
    float2 acc[CH];
#pragma unroll
    for (int i = 0; i < CH; i++) acc[i] = make_float2(threadIdx.x * 1e-3f + i, i * 0.5f);
    const float2 bb = make_float2(b, b * 1.0001f), cc = make_float2(c, c * 0.9999f);
    float2 coef2[CH];
#pragma unroll
    for (int i = 0; i < CH; i++) coef2[i] = make_float2(coef[(threadIdx.x + i) & 3] * 1e-3f, coef[(threadIdx.x + 2 * i) & 3] * 1e-3f);
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < CH; i++) {
            if (MODE == 0) { acc[i].x = fmaf(acc[i].x, bb.x, cc.x); }
            else if (MODE == 1) acc[i] = __ffma2_rn(acc[i], bb, cc);
            else if (MODE == 2) { acc[i].x = fmaf(coef2[i].x, bb.x, acc[i].x); }
            else acc[i] = __ffma2_rn(coef2[i], bb, acc[i]);
        }
    }
    const long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < CH; i++) s += acc[i].x + acc[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int MODE>
static void run(const char* name, int warps) {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float* out; long long* cyc;
    cudaMalloc(&out, sizeof(float) * sms * warps * 32);
    cudaMalloc(&cyc, sizeof(long long) * sms);
    float hc[4] = {0.999f, 0.9991f, 1e-3f, 1.1e-3f};
    float* coef;
    cudaMalloc(&coef, sizeof(hc));
    cudaMemcpy(coef, hc, sizeof(hc), cudaMemcpyHostToDevice);
    for (int rep = 0; rep < 3; rep++) rate<MODE><<<sms, warps * 32>>>(out, cyc, coef);
    cudaDeviceSynchronize();
    long long* h = new long long[sms];
    cudaMemcpy(h, cyc, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
    double mean = 0;
    for (int i = 0; i < sms; i++) mean += double(h[i]) / sms;
    const double inst_per_smsp = double(ITERS) * CH * warps / 4.0;
    const double fma_per_clk_sm = double(ITERS) * CH * warps * 32 * ((MODE & 1) ? 2 : 1) / mean;
    printf("{\"instr\": \"%s\", \"warps_per_sm\": %d, \"cycles\": %.0f, \"cycles_per_instr_per_smsp\": %.3f, \"fma_per_clk_per_sm\": %.1f, "
           "\"err\": \"%s\"}\n", name, warps, mean, mean / inst_per_smsp, fma_per_clk_sm, cudaGetErrorString(cudaGetLastError()));
    cudaFree(out); cudaFree(cyc); cudaFree(coef); delete[] h;
}

int main() {
    for (int w : {4, 8, 16, 32}) {
        run<0>("FFMA acc=acc*b+c", w); run<1>("FFMA2 acc=acc*b+c", w);
        run<2>("FFMA acc=coef_i*x+acc (matvec-like)", w); run<3>("FFMA2 acc=coef_i*x+acc (matvec-like)", w);
    }
    return 0;
}
