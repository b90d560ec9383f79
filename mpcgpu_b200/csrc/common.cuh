// common.cuh -- shared helpers for the sm_100a KKT linear-system kernels.
//
// Thread mapping used everywhere in this directory ("row-owner" layout):
//   one group of LPK=16 lanes per knot; lane i<STATE owns ROW i of every STATE x STATE
//   block of that knot in registers.  Two knots share a warp, so all intra-knot
//   synchronisation is __syncwarp(); cross-knot traffic goes through padded shared-memory
//   tiles (row stride 16 floats = 64 B so a row is four LDS.128) and, across the CTAs of a
//   thread-block cluster, through DSMEM.
#pragma once
#include <cuda_runtime.h>
#include <cooperative_groups.h>
#include <cstdint>

namespace mpcb200 {

namespace cg = cooperative_groups;

constexpr int LPK = 16;  // lanes per knot (STATE <= 16)
constexpr int LD = 16;   // padded leading dimension of shared-memory tiles (floats)

// ---- cluster helpers (no-ops for CL == 1) ---------------------------------------------
template <int CL>
__device__ __forceinline__ void cluster_sync_all() {
    if constexpr (CL == 1) {
        __syncthreads();
    } else {
        cg::this_cluster().sync();
    }
}

template <int CL>
__device__ __forceinline__ unsigned cluster_rank() {
    if constexpr (CL == 1) return 0u;
    else return cg::this_cluster().block_rank();
}

template <int CL, typename P>
__device__ __forceinline__ P* map_rank(P* p, unsigned rank) {
    if constexpr (CL == 1) return p;
    else return cg::this_cluster().map_shared_rank(p, rank);
}

// ---- mbarrier / st.async primitives (cluster kernels): a remote 4-byte store that completes
// transaction bytes on an mbarrier in the DESTINATION CTA.  One-way DSMEM latency, no
// cluster-wide barrier and no membar; the waiter observes the data once the phase flips.
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* mb, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(mb)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init_cluster() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* mb, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(mb)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* mb, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra DONE;\n"
        "bra LAB_WAIT;\n"
        "DONE:\n"
        "}" ::"r"(smem_u32(mb)), "r"(parity) : "memory");
}
__device__ __forceinline__ uint32_t mapa_u32(uint32_t cta_addr, unsigned rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(cta_addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void st_async_f32(uint32_t cluster_addr, float v, uint32_t cluster_mbar) {
    asm volatile("st.async.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(cluster_addr),
                 "r"(__float_as_uint(v)), "r"(cluster_mbar) : "memory");
}

__device__ __forceinline__ void st_async_f64(uint32_t cluster_addr, double v, uint32_t cluster_mbar) {
    asm volatile("st.async.shared::cluster.mbarrier::complete_tx::bytes.b64 [%0], %1, [%2];" ::"r"(cluster_addr),
                 "l"(__double_as_longlong(v)), "r"(cluster_mbar) : "memory");
}
__device__ __forceinline__ void st_async(uint32_t a, float v, uint32_t mb) { st_async_f32(a, v, mb); }
__device__ __forceinline__ void st_async(uint32_t a, double v, uint32_t mb) { st_async_f64(a, v, mb); }

// ---- tensor memory (TMEM, 128 lanes x 512 columns x 32 bit per SM) as an operand store ----------
// tcgen05.ld/st .32x32b: warp w of the CTA addresses lanes 32*(w%4)..+31, thread = lane, xN = N
// consecutive 32-bit columns of that lane.  Measured on B200 (tools/tmem_bw.cu): >= 300 B/clk/SM,
// concurrent with (not taken from) the 128 B/clk shared-memory pipe.
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_word, uint32_t ncols) {   // whole warp; ncols pow2 >= 32
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_word)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {      // whole warp
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

template <int N> struct TmemChunk;   // N floats of one lane: asynchronous load, wait, store
template <> struct TmemChunk<16> {
    uint32_t q[16];
    __device__ __forceinline__ void ld(uint32_t a) {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                     : "=r"(q[0]), "=r"(q[1]), "=r"(q[2]), "=r"(q[3]), "=r"(q[4]), "=r"(q[5]), "=r"(q[6]), "=r"(q[7]),
                       "=r"(q[8]), "=r"(q[9]), "=r"(q[10]), "=r"(q[11]), "=r"(q[12]), "=r"(q[13]), "=r"(q[14]), "=r"(q[15])
                     : "r"(a) : "memory");
    }
    // the registers pass THROUGH the wait so no use of them can be scheduled above it
    __device__ __forceinline__ void wait() {
        asm volatile("tcgen05.wait::ld.sync.aligned;"
                     : "+r"(q[0]), "+r"(q[1]), "+r"(q[2]), "+r"(q[3]), "+r"(q[4]), "+r"(q[5]), "+r"(q[6]), "+r"(q[7]),
                       "+r"(q[8]), "+r"(q[9]), "+r"(q[10]), "+r"(q[11]), "+r"(q[12]), "+r"(q[13]), "+r"(q[14]), "+r"(q[15])
                     :: "memory");
    }
    __device__ __forceinline__ void st(uint32_t a) const {
        asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
                     ::"r"(a), "r"(q[0]), "r"(q[1]), "r"(q[2]), "r"(q[3]), "r"(q[4]), "r"(q[5]), "r"(q[6]), "r"(q[7]),
                       "r"(q[8]), "r"(q[9]), "r"(q[10]), "r"(q[11]), "r"(q[12]), "r"(q[13]), "r"(q[14]), "r"(q[15]) : "memory");
    }
};
template <> struct TmemChunk<8> {
    uint32_t q[8];
    __device__ __forceinline__ void ld(uint32_t a) {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(q[0]), "=r"(q[1]), "=r"(q[2]), "=r"(q[3]), "=r"(q[4]), "=r"(q[5]), "=r"(q[6]), "=r"(q[7])
                     : "r"(a) : "memory");
    }
    __device__ __forceinline__ void wait() {
        asm volatile("tcgen05.wait::ld.sync.aligned;"
                     : "+r"(q[0]), "+r"(q[1]), "+r"(q[2]), "+r"(q[3]), "+r"(q[4]), "+r"(q[5]), "+r"(q[6]), "+r"(q[7]) :: "memory");
    }
    __device__ __forceinline__ void st(uint32_t a) const {
        asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                     ::"r"(a), "r"(q[0]), "r"(q[1]), "r"(q[2]), "r"(q[3]), "r"(q[4]), "r"(q[5]), "r"(q[6]), "r"(q[7]) : "memory");
    }
};
template <> struct TmemChunk<4> {
    uint32_t q[4];
    __device__ __forceinline__ void ld(uint32_t a) {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(q[0]), "=r"(q[1]), "=r"(q[2]), "=r"(q[3]) : "r"(a) : "memory");
    }
    __device__ __forceinline__ void wait() {
        asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(q[0]), "+r"(q[1]), "+r"(q[2]), "+r"(q[3]) :: "memory");
    }
    __device__ __forceinline__ void st(uint32_t a) const {
        asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1,%2,%3,%4};"
                     ::"r"(a), "r"(q[0]), "r"(q[1]), "r"(q[2]), "r"(q[3]) : "memory");
    }
};

// Sum over the warp, the same bits in every lane.  A radix-8 and a radix-4 exchange (7 + 3 INDEPENDENT shuffles) instead of
// the five dependent levels of a butterfly: two shuffle latencies on the critical path instead of five (the PCG loops are
// latency chains: measured ~29 cycles per level).  The additions are those of the butterfly over lane bits 0,1,2 | 3,4.
template <class T>
__device__ __forceinline__ T warp_sum_radix(T a) {
    T s[8];
    s[0] = a;
#pragma unroll
    for (int j = 1; j < 8; j++) s[j] = __shfl_xor_sync(0xffffffffu, a, j);
    const T b = ((s[0] + s[1]) + (s[2] + s[3])) + ((s[4] + s[5]) + (s[6] + s[7]));
    const T u1 = __shfl_xor_sync(0xffffffffu, b, 8), u2 = __shfl_xor_sync(0xffffffffu, b, 16), u3 = __shfl_xor_sync(0xffffffffu, b, 24);
    return (b + u1) + (u2 + u3);
}
#ifndef MPCB200_WARP_SUM_BUTTERFLY
#define MPCB200_WARP_SUM_BUTTERFLY 0      // 1: the five-level butterfly (A/B builds)
#endif
template <class T>
__device__ __forceinline__ T warp_sum_butterfly(T v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_sum(float v) { return MPCB200_WARP_SUM_BUTTERFLY ? warp_sum_butterfly(v) : warp_sum_radix(v); }
__device__ __forceinline__ double warp_sum(double v) { return MPCB200_WARP_SUM_BUTTERFLY ? warp_sum_butterfly(v) : warp_sum_radix(v); }

// ---- row tiles ------------------------------------------------------------------------
// write my row (D values, zero padded to a multiple of 4) to dst + row*ld
template <int D>
__device__ __forceinline__ void publish_row(float* dst, int ld, int row, const float (&a)[D]) {
    float4* d = reinterpret_cast<float4*>(dst + row * ld);
#pragma unroll
    for (int q = 0; q < (D + 3) / 4; q++) {
        float4 v;
        v.x = (4 * q + 0 < D) ? a[4 * q + 0] : 0.f;
        v.y = (4 * q + 1 < D) ? a[4 * q + 1] : 0.f;
        v.z = (4 * q + 2 < D) ? a[4 * q + 2] : 0.f;
        v.w = (4 * q + 3 < D) ? a[4 * q + 3] : 0.f;
        d[q] = v;
    }
}

// read one padded row (broadcast within the knot group) into registers
template <int D>
__device__ __forceinline__ void load_row(float (&r)[D], const float* src) {
    const float4* s = reinterpret_cast<const float4*>(src);
#pragma unroll
    for (int q = 0; q < (D + 3) / 4; q++) {
        float4 v = s[q];
        if (4 * q + 0 < D) r[4 * q + 0] = v.x;
        if (4 * q + 1 < D) r[4 * q + 1] = v.y;
        if (4 * q + 2 < D) r[4 * q + 2] = v.z;
        if (4 * q + 3 < D) r[4 * q + 3] = v.w;
    }
}

// c[j] += sum_m a[m] * Bs[m*ld + j]   (rows of B broadcast from shared memory)
template <int K, int NOUT>
__device__ __forceinline__ void row_gemm_acc(float (&c)[NOUT], const float (&a)[K],
                                             const float* __restrict__ Bs, int ld) {
#pragma unroll
    for (int m = 0; m < K; m++) {
        float b[NOUT];
        load_row<NOUT>(b, Bs + m * ld);
#pragma unroll
        for (int j = 0; j < NOUT; j++) c[j] = fmaf(a[m], b[j], c[j]);
    }
}

// dot of my row with a broadcast vector in shared memory
template <int K>
__device__ __forceinline__ float row_dot(const float (&a)[K], const float* __restrict__ v) {
    float x[K];
    load_row<K>(x, v);
    float acc0 = 0.f, acc1 = 0.f;
#pragma unroll
    for (int m = 0; m + 1 < K; m += 2) {
        acc0 = fmaf(a[m], x[m], acc0);
        acc1 = fmaf(a[m + 1], x[m + 1], acc1);
    }
    if (K & 1) acc0 = fmaf(a[K - 1], x[K - 1], acc0);
    return acc0 + acc1;
}

// In-place Gauss-Jordan inverse WITHOUT pivoting of a D x D matrix held one row per lane
// (lane `i` owns row i; lanes i >= D carry zeros and stay harmless).  Same elimination
// sequence as the reference's invertMatrix on [A | I] (include/utils/matrix.cuh:120-238):
// the in-place form keeps the evolving inverse column in the slot of the eliminated one.
// `scratch`: LD floats per knot group, used to broadcast the pivot row.
template <int D>
__device__ __forceinline__ void gj_invert_rows(float (&a)[D], int i, float* scratch) {
#pragma unroll
    for (int p = 0; p < D; p++) {
        if (i == p) publish_row<D>(scratch, 0, 0, a);
        __syncwarp();
        float rp[D];
        load_row<D>(rp, scratch);
        __syncwarp();
        const float pinv = 1.0f / rp[p];
        const float f = a[p] * pinv;
        if (i == p) {
#pragma unroll
            for (int j = 0; j < D; j++) a[j] = rp[j] * pinv;
            a[p] = pinv;
        } else {
#pragma unroll
            for (int j = 0; j < D; j++)
                if (j != p) a[j] = fmaf(-f, rp[j], a[j]);
            a[p] = -f;
        }
    }
}

}  // namespace mpcb200
