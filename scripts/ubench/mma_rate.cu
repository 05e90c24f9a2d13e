// Micro-benchmark: rate of tcgen05.mma (cta_group::1, kind::f16, M = 128, K = 16) with BOTH operands in shared
// memory in the no-swizzle K-major layout of klm_seq_tc, as a function of N, alone and with bulk copies landing in
// the same SM's shared memory at the same time.  One CTA per SM, all SMs busy.
#include <cstdio>
#include <cuda_runtime.h>
#include "../../ocrd_keraslm_b200/csrc/klm_tc.cuh"

#define W 512
#define KC 64
#define CHUNK (KC * 256)
#define RING 4

__global__ void __launch_bounds__(128, 1) k(const unsigned char *g, long long *out, int N2, int copies, int reps) {
    extern __shared__ __align__(1024) unsigned char smem[];
    unsigned char *ring = smem;                        // RING chunks of A
    unsigned char *land = smem + RING * CHUNK;         // landing zone of the concurrent copies (1 chunk)
    unsigned char *wsm = land + CHUNK;                 // B image [W/8][N2][8] halves
    uint64_t *bars = reinterpret_cast<uint64_t *>(wsm + (size_t)W * N2 * 2);
    uint32_t *slot = reinterpret_cast<uint32_t *>(bars + 4);
    const uint32_t bar_done = tc::smem_u32(bars), bar_cp = tc::smem_u32(bars + 1);
    const int warp = threadIdx.x >> 5;
    if (warp == 0) { tc::tmem_alloc(tc::smem_u32(slot), 256); tc::tmem_relinquish(); }
    if (threadIdx.x == 32) { tc::mbar_init(bar_done, 1); tc::mbar_init(bar_cp, 1); tc::fence_mbar_init(); }
    for (int i = threadIdx.x; i < (RING * CHUNK + CHUNK + W * N2 * 2) / 4; i += blockDim.x) reinterpret_cast<unsigned *>(smem)[i] = 0;
    tc::fence_proxy_async();
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem = *slot;
    if (threadIdx.x == 0) {
        const uint32_t idesc = tc::idesc_f16(128, N2);
        const uint32_t lbo_a = 128 * 16, lbo_b = N2 * 16;
        const uint32_t ring_a = tc::smem_u32(ring), w_a = tc::smem_u32(wsm);
        long long total = 0;
        for (int r = 0; r < reps; ++r) {
            const long long t0 = clock64();
            for (int c = 0; c < W / KC; ++c) {
#pragma unroll
                for (int kk = 0; kk < KC / 16; ++kk) {
                    const uint64_t ad = tc::smem_desc(ring_a + (c % RING) * CHUNK + kk * 2 * lbo_a, lbo_a, 128);
                    const uint64_t bd = tc::smem_desc(w_a + (uint32_t)(c * (KC / 8) + kk * 2) * lbo_b, lbo_b, 128);
                    tc::mma_f16(tmem, ad, bd, idesc, (c | kk) != 0 ? 1u : 0u);
                }
            }
            tc::mma_commit(bar_done);
            while (!tc::mbar_try_wait(bar_done, r & 1)) { }
            total += clock64() - t0;
        }
        out[blockIdx.x] = total / reps;
    } else if (threadIdx.x == 64 && copies) {
        // keep `copies` chunk-sized bulk copies per MMA pass landing in this SM's shared memory
        for (int r = 0; r < reps; ++r) {
            tc::mbar_arrive_expect_tx(bar_cp, (uint32_t)copies * CHUNK);
            for (int c = 0; c < copies; ++c)
                tc::bulk_g2s(tc::smem_u32(land), g + (size_t)((blockIdx.x + c) % 8) * CHUNK, CHUNK, bar_cp);
            while (!tc::mbar_try_wait(bar_cp, r & 1)) { }
        }
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 0) { __syncwarp(); tc::tmem_dealloc(tmem, 256); }
}

int main() {
    unsigned char *g; long long *out;
    cudaMalloc(&g, 8 * CHUNK); cudaMemset(g, 0, 8 * CHUNK);
    cudaMalloc(&out, 148 * 8);
    for (int copies : {0, 8})
        for (int N2 : {64, 96, 128, 160, 192, 256}) {
            const size_t smem = RING * CHUNK + CHUNK + (size_t)W * N2 * 2 + 128;
            if (smem > 232448) { printf("N %3d: B image does not fit\n", N2); continue; }
            cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            k<<<148, 128, smem>>>(g, out, N2, copies, 200);
            cudaError_t e = cudaDeviceSynchronize();
            long long h[148];
            cudaMemcpy(h, out, sizeof h, cudaMemcpyDeviceToHost);
            double mean = 0; for (int i = 0; i < 148; ++i) mean += h[i]; mean /= 148;
            printf("N %3d, %d x 16 KB bulk copies in flight per pass: %6.0f clk per 32 MMAs = %5.1f clk per MMA (floor N/2 = %d)  (%s)\n",
                   N2, copies, mean, mean / 32, N2 / 2, cudaGetErrorString(e));
        }
    return 0;
}
