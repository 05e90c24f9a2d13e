// Micro-benchmark: how fast can one SM fill its shared memory?  (a) bulk copy from L2, (b) bulk copy from the
// peer CTA's shared memory (DSMEM), (c) both at once, (d) L2 bulk copy multicast to both CTAs of a pair.
// Every CTA pair of the grid does the same thing at the same time (all SMs busy), SM clocks of CTA 0 are printed.
#include <cstdio>
#include <cuda_runtime.h>
#include "../../ocrd_keraslm_b200/csrc/klm_tc.cuh"

#define CHUNK 16384
#define NCH 4

__device__ __forceinline__ void bulk_s2s(uint32_t dst_cluster, uint32_t src_cta, uint32_t bytes, uint32_t bar_cluster) {
    asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst_cluster), "r"(src_cta), "r"(bytes), "r"(bar_cluster) : "memory");
}
__device__ __forceinline__ void bulk_g2s_mc(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar, uint16_t mask) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar), "h"(mask) : "memory");
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(256, 1) k(unsigned char *g, long long *out, int mode, int reps) {
    extern __shared__ __align__(1024) unsigned char smem[];
    unsigned char *bufA = smem;                    // NCH chunks: landing zone of the L2 copies
    unsigned char *bufB = smem + NCH * CHUNK;      // NCH chunks: landing zone of the peer's copies
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + 2 * NCH * CHUNK);
    const uint32_t rank = tc::cluster_ctarank();
    const uint32_t barA = tc::smem_u32(bars), barB = tc::smem_u32(bars + 1);
    if (threadIdx.x == 0) { tc::mbar_init(barA, 1); tc::mbar_init(barB, 1); tc::fence_mbar_init(); }
    __syncthreads();
    tc::cluster_sync();
    long long t_total = 0;
    for (int r = 0; r < reps; ++r) {
        uint32_t expectA = 0, expectB = 0;
        if (mode == 0 || mode == 2 || mode == 3 || mode >= 4) expectA = NCH * CHUNK;   // mode 3: half from own multicast, half from the peer's
        if (mode == 1 || mode == 2) expectB = NCH * CHUNK;
        if (threadIdx.x == 0) {
            if (expectA) tc::mbar_arrive_expect_tx(barA, expectA);
            if (expectB) tc::mbar_arrive_expect_tx(barB, expectB);
        }
        if (mode >= 4) {
            // the PEER's source region is (re)written first by this CTA's generic stores (fresh lines, other SM)
            unsigned char *dstw = g + ((size_t)((blockIdx.x >> 1) * 7 + r) % 64) * (NCH * CHUNK);
            if (mode == 4) {          // scattered 4-byte stores, like the kernel's hidden-vector publication
                unsigned *w = reinterpret_cast<unsigned *>(dstw) + (size_t)rank * (NCH * CHUNK / 8);
                for (int i = threadIdx.x; i < NCH * CHUNK / 8; i += blockDim.x) w[(i * 37) % (NCH * CHUNK / 8)] = i + r;
            } else {                  // 16-byte coalesced stores
                uint4 *w = reinterpret_cast<uint4 *>(dstw) + (size_t)rank * (NCH * CHUNK / 32);
                for (int i = threadIdx.x; i < NCH * CHUNK / 32; i += blockDim.x) w[i] = make_uint4(i, r, i, r);
            }
            __threadfence();
            tc::fence_proxy_async();
        }
        tc::cluster_sync();                            // both CTAs have armed their barriers
        if (threadIdx.x == 0) {
            const long long t0 = clock64();
            const unsigned char *src = g + ((size_t)((blockIdx.x >> 1) * 7 + r) % 64) * (NCH * CHUNK);
            for (int c = 0; c < NCH; ++c) {
                if (mode == 0 || mode == 2 || mode >= 4) tc::bulk_g2s(tc::smem_u32(bufA + c * CHUNK), src + c * CHUNK, CHUNK, barA);
                if (mode == 1 || mode == 2)
                    bulk_s2s(tc::mapa(tc::smem_u32(bufB + c * CHUNK), rank ^ 1), tc::smem_u32(bufA + c * CHUNK), CHUNK,
                             tc::mapa(barB, rank ^ 1));
                if (mode == 3 && (c & 1) == (int)rank)
                    bulk_g2s_mc(tc::smem_u32(bufA + c * CHUNK), src + c * CHUNK, CHUNK, barA, 3);
            }
            if (expectA) while (!tc::mbar_try_wait(barA, r & 1)) { }
            if (expectB) while (!tc::mbar_try_wait(barB, r & 1)) { }
            t_total += clock64() - t0;
        }
        __syncthreads();
    }
    tc::cluster_sync();
    if (threadIdx.x == 0) out[blockIdx.x] = t_total / reps;
}

int main() {
    unsigned char *g; long long *out;
    cudaMalloc(&g, 64 * NCH * CHUNK); cudaMemset(g, 1, 64 * NCH * CHUNK);
    cudaMalloc(&out, 148 * 8);
    const size_t smem = 2 * NCH * CHUNK + 64;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const char *names[] = {"L2 -> smem (64 KB per SM)", "peer smem -> smem (64 KB per SM)", "both at once (64 + 64 KB)",
                           "L2 multicast to the pair (64 KB per SM, 32 KB requested)",
                           "L2 -> smem, source just rewritten by 4-byte stores", "L2 -> smem, source just rewritten by 16-byte stores"};
    for (int grid : {2, 148})
        for (int mode = 0; mode < 6; ++mode) {
            k<<<grid, 256, smem>>>(g, out, mode, 50);
            cudaError_t e = cudaDeviceSynchronize();
            long long h[148];
            cudaMemcpy(h, out, grid * 8, cudaMemcpyDeviceToHost);
            double mean = 0; for (int i = 0; i < grid; ++i) mean += h[i]; mean /= grid;
            printf("grid %3d  %-58s : %8.0f clk  -> %.1f B/clk per SM  (%s)\n", grid, names[mode], mean,
                   (mode == 2 ? 2.0 : 1.0) * NCH * CHUNK / mean, cudaGetErrorString(e));
        }
    return 0;
}
