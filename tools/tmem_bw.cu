// tmem_bw.cu -- microbenchmark: can tensor memory (TMEM, 256 KB/SM) serve as a second operand store
// next to shared memory for the PCG matvec?  Measures tcgen05.ld (32x32b.x32) read throughput per SM
// with 4 and 8 warps, alone and concurrently with LDS.128 traffic, in bytes per clock per SM.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 tools/tmem_bw.cu -o tmem_bw
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
          "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]),
          "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]),
          "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31]) : "memory");
}

// mode bit0: TMEM loads, bit1: LDS.128 loads (4 wavefronts per warp instruction, conflict-free)
__global__ void __launch_bounds__(256) k(int mode, int reps, unsigned long long* cyc, float* sink) {
    __shared__ uint32_t tbase;
    extern __shared__ float4 sm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 40 * 256; i += blockDim.x) sm[i] = make_float4(i, 1, 2, 3);
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(&tbase)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tb = tbase;
    // warp w owns lanes 32*(w%4).., columns 256*(w/4) .. +255
    const uint32_t mine = tb + ((uint32_t)(32 * (warp & 3)) << 16) + 256u * (warp >> 2);
    uint32_t r[32];
    for (int i = 0; i < 32; i++) r[i] = __float_as_uint((float)(threadIdx.x + i));
    for (int c = 0; c < 256; c += 32) tmem_st32(mine + c, r);
    asm volatile("tcgen05.wait::st.sync.aligned;");
    __syncthreads();
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    const unsigned long long t0 = clock64();
    for (int it = 0; it < reps; it++) {
        if (mode & 1) {
            uint32_t q[5][32];
#pragma unroll
            for (int c = 0; c < 5; c++) tmem_ld32(mine + 32 * c, q[c]);   // 160 floats per thread (a Pinv-sized read)
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int c = 0; c < 5; c++)
#pragma unroll
                for (int i = 0; i < 32; i++) acc[i & 7] += __uint_as_float(q[c][i]);
        }
        if (mode & 2) {
#pragma unroll
            for (int c = 0; c < 40; c++) {           // 40 x LDS.128 = 160 floats per thread, all distinct
                const float4 v = sm[c * 256 + ((threadIdx.x + it) & 255)];
                acc[c & 7] += (v.x + v.y) + (v.z + v.w);
            }
        }
    }
    const unsigned long long t1 = clock64();
    __syncthreads();
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    sink[blockIdx.x * blockDim.x + threadIdx.x] = ((acc[0] + acc[1]) + (acc[2] + acc[3])) + ((acc[4] + acc[5]) + (acc[6] + acc[7]));
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tb));
}

int main() {
    unsigned long long* cyc; float* sink;
    cudaMalloc(&cyc, 148 * 8); cudaMalloc(&sink, 148 * 256 * 4);
    const int reps = 2000;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 40 * 256 * 16);
    for (int threads : {128, 256}) for (int mode : {1, 2, 3}) {
        k<<<148, threads, 40 * 256 * 16>>>(mode, 10, cyc, sink);
        k<<<148, threads, 40 * 256 * 16>>>(mode, reps, cyc, sink);
        unsigned long long h[148];
        cudaError_t e = cudaMemcpy(h, cyc, sizeof h, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
        const double bytes = (double)reps * threads * 160 * 4 * ((mode == 3) ? 2 : 1);
        printf("{\"threads\": %d, \"mode\": \"%s\", \"cycles\": %llu, \"bytes_per_clk_per_sm\": %.1f}\n", threads,
               mode == 1 ? "tmem" : mode == 2 ? "lds128" : "tmem+lds128", h[0], bytes / (double)h[0]);
    }
    return 0;
}
