// occupancy_probe.cu -- what the device reports for the L1-shim pcg<float,14,N> kernels behind the
// reference's launch (cooperative, grid = N, 128 threads, include/pcg/sqp.cuh:275): registers, shared
// memory, CTAs per SM, co-resident clusters.  Build (from the repo root):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I mpcgpu_b200/shim -I include tools/occupancy_probe.cu -o /tmp/occupancy_probe
#include <cstdio>
#include "gpu_pcg.cuh"

template <uint32_t N>
static void probe() {
    void* k = (void*)pcg<float, 14, N>;
    const size_t smem = pcgSharedMemSize<float>(14, N);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const unsigned cl = mpcb200_shim::Use<float, 14, N>::DIM;
    if (cl > 8) cudaFuncSetAttribute(k, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, k);
    int per_sm = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, 128, smem);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(N); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = cl; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeCooperative; at[1].val.cooperative = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    int ncl = -1, ncl_coop = -1;
    cudaError_t e1 = cudaOccupancyMaxActiveClusters(&ncl, k, &cfg);
    cfg.numAttrs = 2;
    cudaError_t e2 = cudaOccupancyMaxActiveClusters(&ncl_coop, k, &cfg);
    printf("{\"N\": %u, \"pinv_in_tmem\": %d, \"cluster\": %u, \"regs\": %d, \"smem\": %zu, \"ctas_per_sm\": %d, \"max_clusters\": %d, \"max_clusters_coop\": %d, \"err\": [%d, %d], \"needed_clusters\": %u}\n",
           N, mpcb200_shim::Use<float, 14, N>::CLUSTER ? (int)mpcb200_shim::Tiling<float, N>::TM : -1, cl, fa.numRegs, smem, per_sm, ncl, ncl_coop, (int)e1, (int)e2, N / cl);
    cudaGetLastError();
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    printf("{\"device\": \"%s\", \"sms\": %d, \"smem_per_sm\": %zu, \"regs_per_sm\": %d}\n", p.name, p.multiProcessorCount, p.sharedMemPerMultiprocessor, p.regsPerMultiprocessor);
    probe<8>(); probe<32>(); probe<64>(); probe<128>(); probe<256>(); probe<512>();
    return 0;
}
