// Micro-benchmark: SM-to-SM signalling latency through L2.  CTA 0 plays ping-pong with CTA k (k = 1..grid-1, one
// after the other): it stores a sequence number into k's mailbox, k polls, answers into 0's mailbox, 0 polls.
// Reported: clocks per round trip / 2 for every partner (their SM id shows the die / GPC), for
//   mode 0: st.relaxed.gpu + ld.relaxed.gpu polling     mode 1: st.volatile + ld.volatile
//   mode 2: red.release.gpu (atomic add) + ld.acquire.gpu polling
//   mode 3: like 0, but 8-byte {value, seq} words and 32 lanes polling 32 words (the exchange of klm_seq_fma_ll)
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned smid() { unsigned r; asm volatile("mov.u32 %0, %%smid;" : "=r"(r)); return r; }
__device__ __forceinline__ void st_rel(unsigned *p, unsigned v, int mode) {
    if (mode == 1) asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
    else if (mode == 2) asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(1u) : "memory");
    else asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acq(const unsigned *p, int mode) {
    unsigned v;
    if (mode == 1) asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    else if (mode == 2) asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    else asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// mailboxes: 256 bytes per CTA
__global__ void k(unsigned *mbox, long long *out, unsigned *sm, int mode, int reps) {
    const int me = blockIdx.x, G = gridDim.x;
    if (threadIdx.x == 0) sm[me] = smid();
    if (mode == 3) {
        uint2 *mb2 = reinterpret_cast<uint2 *>(mbox);
        const int lane = threadIdx.x;
        if (me == 0) {
            unsigned seq = 0;
            for (int kk = 1; kk < G; ++kk) {
                long long t0 = clock64();
                for (int r = 0; r < reps; ++r) {
                    ++seq;
                    asm volatile("st.relaxed.gpu.global.v2.u32 [%0], {%1, %2};" ::"l"(mb2 + kk * 32 + lane), "r"(lane), "r"(seq) : "memory");
                    uint2 v;
                    do { asm volatile("ld.relaxed.gpu.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(mb2 + lane) : "memory"); } while (v.y != seq);
                    __syncwarp();
                }
                if (lane == 0) out[kk] = (clock64() - t0) / reps / 2;
            }
            // release everybody
            asm volatile("st.relaxed.gpu.global.v2.u32 [%0], {%1, %2};" ::"l"(mb2 + G * 32 + lane), "r"(1u), "r"(1u) : "memory");
        } else {
            unsigned seq = (unsigned)(me - 1) * reps;
            for (int r = 0; r < reps; ++r) {
                ++seq;
                uint2 v;
                do { asm volatile("ld.relaxed.gpu.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(mb2 + me * 32 + lane) : "memory"); } while (v.y != seq);
                __syncwarp();
                asm volatile("st.relaxed.gpu.global.v2.u32 [%0], {%1, %2};" ::"l"(mb2 + lane), "r"(lane), "r"(seq) : "memory");
            }
        }
        return;
    }
    if (threadIdx.x != 0) return;
    unsigned *mine = mbox + me * 64, *zero = mbox;
    if (me == 0) {
        unsigned seq = 0;
        for (int kk = 1; kk < G; ++kk) {
            unsigned *their = mbox + kk * 64;
            long long t0 = clock64();
            for (int r = 0; r < reps; ++r) {
                ++seq;
                st_rel(their, seq, mode);
                while (ld_acq(zero, mode) != seq) { }
            }
            out[kk] = (clock64() - t0) / reps / 2;
        }
    } else {
        unsigned seq = (unsigned)(me - 1) * reps, mine_seq = 0;
        for (int r = 0; r < reps; ++r) {
            ++seq; ++mine_seq;
            while (ld_acq(mine, mode) != (mode == 2 ? mine_seq : seq)) { }
            st_rel(zero, seq, mode);
        }
    }
}

int main() {
    unsigned *mbox, *sm; long long *out;
    cudaMalloc(&mbox, 149 * 256); cudaMalloc(&sm, 148 * 4); cudaMalloc(&out, 148 * 8);
    const char *names[] = {"st/ld.relaxed.gpu", "st/ld.volatile", "red.release + ld.acquire", "32 x {value,seq} words, relaxed.gpu"};
    for (int mode = 0; mode < 4; ++mode) {
        cudaMemset(mbox, 0, 149 * 256);
        int reps = 200;
        void *args[] = {&mbox, &out, &sm, &mode, &reps};
        cudaLaunchCooperativeKernel((const void *)k, dim3(148), dim3(32), args, 0, 0);
        cudaError_t e = cudaDeviceSynchronize();
        long long h[148]; unsigned s[148];
        cudaMemcpy(h, out, sizeof h, cudaMemcpyDeviceToHost); cudaMemcpy(s, sm, sizeof s, cudaMemcpyDeviceToHost);
        long long mn = 1 << 30, mx = 0; double mean = 0;
        for (int i = 1; i < 148; ++i) { mn = h[i] < mn ? h[i] : mn; mx = h[i] > mx ? h[i] : mx; mean += h[i]; }
        printf("mode %d (%s): one-way latency min %lld mean %.0f max %lld clk  (%s); CTA 0 on SM %u\n", mode, names[mode], mn, mean / 147, mx,
               cudaGetErrorString(e), s[0]);
        if (mode == 0 || mode == 3) {
            printf("  by partner SM id:");
            for (int i = 1; i < 148; ++i) printf(" %u:%lld", s[i], h[i]);
            printf("\n");
        }
    }
    return 0;
}
