// Micro-benchmark of the hidden-vector hand-over of klm_seq_tc: every tick all CTAs of the grid publish a slice of
// one 128 KB block and, after a grid barrier, every CTA pulls the WHOLE block into its shared memory with bulk
// copies.  Question: what limits the pull (measured ~17-20 B/clk per SM inside the kernel, 57 B/clk in isolation)?
//   write mode: 0 = block never rewritten, 1 = each CTA rewrites its slice with coalesced 16-byte stores,
//               2 = each CTA rewrites its slice with scattered 4-byte stores, 3 = slice rewritten by one bulk store
//   read mode:  0 = all CTAs read the same block, chunk order rotated by CTA, 1 = same block, same order,
//               2 = every CTA reads a private block (no sharing), 3 = same block through NREP replicas (CTA % 4),
//               4 = same block, ld.global.cg into registers + st.shared (generic path, 256 threads)
#include <cstdio>
#include <cuda_runtime.h>
#include "../../ocrd_keraslm_b200/csrc/klm_tc.cuh"

#define BLOCK_BYTES (128 * 1024)
#define CHUNK 16384
#define NCH (BLOCK_BYTES / CHUNK)
#define NREP 4

__device__ __forceinline__ void grid_barrier(unsigned *ctr, unsigned target) {
    __syncthreads();
    if (threadIdx.x == 0) {
        tc::red_release_add(ctr, 1u);
        while (tc::ld_acquire(ctr) < target) { }
    }
    __syncthreads();
}

__global__ void __launch_bounds__(256, 1) k(unsigned char *g, unsigned *ctr, long long *out, int wmode, int rmode, int reps,
                                            int fence) {
    extern __shared__ __align__(1024) unsigned char smem[];
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + BLOCK_BYTES);
    unsigned char *stage = smem + BLOCK_BYTES + 64;
    const uint32_t bar = tc::smem_u32(bars);
    const unsigned G = gridDim.x;
    if (threadIdx.x == 0) { tc::mbar_init(bar, 1); tc::fence_mbar_init(); }
    __syncthreads();
    long long t_total = 0, t_max = 0;
    const size_t slice = (BLOCK_BYTES / G) & ~(size_t)15;          // bytes of the block this CTA rewrites
    for (int r = 0; r < reps; ++r) {
        unsigned char *blk = g + (size_t)(r & 1) * BLOCK_BYTES;     // double-buffered like the kernel's hx
        if (wmode == 1) {
            uint4 *w = reinterpret_cast<uint4 *>(blk + blockIdx.x * slice);
            for (int i = threadIdx.x; i < (int)(slice / 16); i += blockDim.x) w[i] = make_uint4(i, r, i, r);
        } else if (wmode == 2) {
            // 4-byte pieces 16 bytes apart, the other pieces of a sector come from three other CTAs
            unsigned *w = reinterpret_cast<unsigned *>(blk) + (blockIdx.x & 3);
            const int n = (int)(slice / 4), base = (blockIdx.x >> 2) * n * 4;
            for (int i = threadIdx.x; i < n; i += blockDim.x) w[(size_t)(base + i) * 4 % (BLOCK_BYTES / 4 - 4)] = i + r;
        } else if (wmode == 3) {
            uint4 *w = reinterpret_cast<uint4 *>(stage);
            for (int i = threadIdx.x; i < (int)(slice / 16); i += blockDim.x) w[i] = make_uint4(i, r, i, r);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncthreads();
            if (threadIdx.x == 0) {
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                             ::"l"(blk + blockIdx.x * slice), "r"(tc::smem_u32(stage)), "r"((uint32_t)slice) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
            }
        }
        if (wmode == 1 || wmode == 2) { __threadfence(); tc::fence_proxy_async(); }
        grid_barrier(ctr, (unsigned)(r + 1) * G);
        const long long t0 = clock64();
        const unsigned char *src = blk;
        if (rmode == 2) src = g + (size_t)(2 + blockIdx.x) * BLOCK_BYTES;
        if (rmode == 3) src = g + (size_t)(2 + (blockIdx.x % NREP)) * BLOCK_BYTES;
        if (rmode == 4) {
            const uint4 *s4 = reinterpret_cast<const uint4 *>(src);
            uint4 *d4 = reinterpret_cast<uint4 *>(smem);
            for (int c0 = 0; c0 < NCH; ++c0) {
                const int c = (c0 + blockIdx.x) % NCH;
                uint4 v[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) v[j] = __ldcg(s4 + c * 1024 + j * 256 + threadIdx.x);
#pragma unroll
                for (int j = 0; j < 4; ++j) d4[c * 1024 + j * 256 + threadIdx.x] = v[j];
            }
            __syncthreads();
        } else if (threadIdx.x == 0) {
            if (fence) tc::fence_proxy_async();
            tc::mbar_arrive_expect_tx(bar, BLOCK_BYTES);
            for (int c0 = 0; c0 < NCH; ++c0) {
                const int c = rmode == 1 ? c0 : (c0 + blockIdx.x) % NCH;
                tc::bulk_g2s(tc::smem_u32(smem + c * CHUNK), src + c * CHUNK, CHUNK, bar);
            }
            while (!tc::mbar_try_wait(bar, r & 1)) { }
        }
        const long long dt = clock64() - t0;
        t_total += dt; t_max = dt > t_max ? dt : t_max;
        __syncthreads();
    }
    if (threadIdx.x == 0) { out[blockIdx.x] = t_total / reps; out[gridDim.x + blockIdx.x] = t_max; }
}

int main(int argc, char **argv) {
    unsigned char *g; unsigned *ctr; long long *out;
    const size_t gbytes = (size_t)(2 + 148) * BLOCK_BYTES;
    cudaMalloc(&g, gbytes); cudaMemset(g, 1, gbytes);
    cudaMalloc(&ctr, 4);
    cudaMalloc(&out, 2 * 148 * 8);
    const size_t smem = BLOCK_BYTES + 64 + 2048;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const char *wn[] = {"no rewrite", "16 B coalesced", "4 B scattered", "bulk store"};
    const char *rn[] = {"same block, rotated", "same block, same order", "private blocks", "4 replicas", "same block, ld.cg+st.shared"};
    for (int grid : {148, 36})
        for (int rmode = 0; rmode < 5; ++rmode)
            for (int wmode = 0; wmode < 4; ++wmode) {
                if (rmode >= 2 && rmode <= 3 && wmode != 0) continue;      // private / replica blocks are never rewritten
                for (int fence = 0; fence < 2; ++fence) {
                    if (rmode == 4 && fence) continue;
                    cudaMemset(ctr, 0, 4);
                    int reps = 200;
                    void *args[] = {&g, &ctr, &out, &wmode, &rmode, &reps, &fence};
                    cudaLaunchCooperativeKernel((const void *)k, dim3(grid), dim3(256), args, smem, 0);
                    cudaError_t e = cudaDeviceSynchronize();
                    long long h[2 * 148];
                    cudaMemcpy(h, out, 2 * 148 * 8, cudaMemcpyDeviceToHost);
                    double mean = 0, worst = 0;
                    for (int i = 0; i < grid; ++i) { mean += h[i]; worst = h[i] > worst ? h[i] : worst; }
                    mean /= grid;
                    printf("grid %3d  read: %-28s write: %-15s fence %d : mean %6.0f clk (%.1f B/clk per SM), slowest CTA %6.0f  (%s)\n",
                           grid, rn[rmode], wn[wmode], fence, mean, BLOCK_BYTES / mean, worst, cudaGetErrorString(e));
                }
            }
    return 0;
}
