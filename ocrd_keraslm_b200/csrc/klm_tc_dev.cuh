// Device helpers shared by the tensor-core sequence kernels (klm_seq_tc.cu, klm_seq_df.cu):
// bounded waits with a device-side watchdog, timeline tracing, the LSTM cell, the LSE stage.
#pragma once
#include "klm_common.cuh"
#include "klm_tc.cuh"

#define TC_SPIN_LIMIT (1u << 24)          // default of KlmTcParams::spin_limit

__device__ __forceinline__ bool tc_spin_fail(unsigned &n, const KlmTcParams &p, int code) {
    if ((++n & 0x3FFu) == 0) {
        if (*(volatile int *)p.err != 0) return true;
        if (n > p.spin_limit) { atomicCAS(p.err, 0, code); return true; }
    }
    return false;
}
// Measured alternatives to the acquire load per poll (KLM_TC_POLL): relaxed polls + one fence.acq_rel.gpu after the
// successful poll are 8 % SLOWER on cfg2 (the fence costs more than it saves), relaxed polls without any fence are
// no faster -- the acquire is not what the barrier costs.
__device__ __forceinline__ bool tc_wait_grid(const KlmTcParams &p, unsigned target, int code) {
    unsigned n = 0;
    if (p.poll_mode == 0) {
        while (tc::ld_acquire(p.barrier) < target)
            if (tc_spin_fail(n, p, code)) return false;
        return true;
    }
    for (;;) {
        unsigned v;
        asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p.barrier) : "memory");
        if (v >= target) break;
        if (tc_spin_fail(n, p, code)) return false;
    }
    if (p.poll_mode == 1) asm volatile("fence.acq_rel.gpu;" ::: "memory");
    return true;
}
__device__ __forceinline__ bool tc_wait_mbar(uint32_t bar, uint32_t parity, const KlmTcParams &p, int code) {
    unsigned n = 0;
    while (!tc::mbar_try_wait(bar, parity))
        if (tc_spin_fail(n, p, code)) return false;
    return true;
}
// optional timeline instrumentation (klm_tc_trace): SM clock of key events for KLM_TC_TRACE_TICKS ticks
__device__ __forceinline__ void tc_trace(const KlmTcParams &p, int tick, int ev) {
    if (p.trace) {
        const int k = tick - p.trace_t0;
        if (k >= 0 && k < KLM_TC_TRACE_TICKS && ev < 8)
        {
            unsigned long long now;
            if (p.trace_mode == 5) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));     // ns, comparable across SMs
            else now = (unsigned long long)clock64();
            p.trace[((size_t)blockIdx.x * KLM_TC_TRACE_TICKS + k) * 8 + ev] = now;
        }
    }
}
__device__ __forceinline__ void tc_bar_epi() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// exp via ex2.approx (2 ulp), reciprocal via rcp.approx: absolute error of the gates ~1e-7
// (ex2.approx.ftz directly: the argument is clamped to +-30 * log2(e), far from the denormal range that exp2f()
//  guards with a compare and two predicated multiplies per call)
__device__ __forceinline__ float tc_exp(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x * 1.4426950408889634f));
    return y;
}
__device__ __forceinline__ float tc_clamp(float x) { return fminf(fmaxf(x, -30.f), 30.f); }

// LSTM cell (Keras LSTMCell, gate order i|f|c|o, exact logistic; rating.py:133-134), 8 MUFU ops:
//   i*g = (1 - e^-2g) / ((1 + e^-i)(1 + e^-2g)),  f = 1/(1 + e^-f),  h = o*tanh(c) likewise
__device__ __forceinline__ void tc_cell(float ai, float af, float ag, float ao, float &c, float &h) {
    const float ei = tc_exp(-tc_clamp(ai)), ef = tc_exp(-tc_clamp(af)), eg = tc_exp(-2.f * tc_clamp(ag)),
                eo = tc_exp(-tc_clamp(ao));
    float num, den;
    klm_tanh_parts(ag, eg, num, den);
    const float ig = __fdividef(num, (1.f + ei) * den);
    const float f = __fdividef(1.f, 1.f + ef);
    c = fmaf(f, c, ig);
    const float ec = tc_exp(-2.f * tc_clamp(c));
    klm_tanh_parts(c, ec, num, den);
    h = __fdividef(num, (1.f + eo) * den);
}

// Two adjacent float4 (32 bytes, 32-byte aligned) of a read-only table with ONE 256-bit load (LDG.E.256): the table-row
// gathers of layer 1 are scattered over the lanes (every stream has its own character), so an instruction costs a sector
// per lane whatever it fetches from it -- a pair of units per instruction halves the sectors touched per value.
__device__ __forceinline__ void tc_ldg_pair(const float4 *p, float4 &a, float4 &b) {
    asm volatile("ld.global.nc.v8.f32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(p));
}

// hidden value -> fp16 hi/lo operand pair, scaled by KLM_TC_HSCALE so that the lo part stays a normal fp16 number
__device__ __forceinline__ void tc_split(float v, __half &hi, __half &lo) {
    v *= KLM_TC_HSCALE;
    hi = __float2half_rn(v);
    lo = __float2half_rn(v - __half2float(hi));
}

// ---------------------------------------------------------------------------------------------
// LSE stage: log-sum-exp over the logits of every stream + pick of the target's logit (or the full softmax row).
// The logits arrive as [parity][Vp/4][64 streams] float4 (the PROJ stores are coalesced over streams).  Lane = stream:
// a warp covers 32 streams and one slice of the vocabulary, reads whole 512-byte rows and keeps a running
// (max, rescaled sum) per lane; the slices of a stream group are combined through shared memory.  One CTA handles
// all 64 streams in about two L2 round trips per tick.
#define TC_LSE_WARPS (KLM_TC_THREADS / 32)
static __device__ void tc_lse_role(const KlmTcParams &p, const KlmTcRole &role, unsigned char *dyn_smem) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const unsigned G = gridDim.x;
    const int nticks = p.T_max + p.skew * p.D;
    const int V = p.V, Vq = p.Vp / 4, V4 = (V + 3) / 4, B = p.B, sigma = p.skew * p.D, full_rows = p.full_rows;
    const int SB = p.wide ? 128 : 64;                      // streams per row of the logits buffer
    const float4 *logits4 = reinterpret_cast<const float4 *>(p.logits);
    const int *targets = p.targets;
    float *out = p.out;
    __shared__ int ok_s;
    // partial (max, sum) per warp and lane: in the dynamic shared memory (an LSE CTA has no other use for it; static
    // shared memory would push the GEMM CTAs' carve-up over the per-block limit)
    float (*part_m)[32] = reinterpret_cast<float (*)[32]>(dyn_smem);
    float (*part_s)[32] = reinterpret_cast<float (*)[32]>(dyn_smem + TC_LSE_WARPS * 32 * sizeof(float));
    // stream groups of 32: this CTA takes groups first_pair, first_pair + npairs, ...; its warps are split evenly
    // over (group, vocabulary slice)
    const int ngroups = (B + 31) / 32;
    const int gmine = role.first_pair < ngroups ? (ngroups - role.first_pair + role.npairs - 1) / role.npairs : 0;
    const int nslices = gmine > 0 ? TC_LSE_WARPS / gmine : 1;
    const int gl = warp / nslices, sl = warp - gl * nslices;              // local group, slice
    const bool warp_on = gmine > 0 && gl < gmine;
    const int s = warp_on ? 32 * (role.first_pair + gl * role.npairs) + lane : 0;
    long long o0 = 0;
    int len = 0;
    if (warp_on && s < B) { o0 = p.offsets[s]; len = (int)(p.offsets[s + 1] - o0); }
    const int v4_lo = (int)((long long)V4 * sl / nslices), v4_hi = (int)((long long)V4 * (sl + 1) / nslices);
    if (tid == 0) tc::red_release_add(p.barrier, 1u);
    for (int tick = 0; tick < nticks; ++tick) {
        if (tid == 0) ok_s = tc_wait_grid(p, (unsigned)(tick + 1) * G, 10) ? 1 : 0;
        __syncthreads();
        if (!ok_s) return;
        if (p.trace && p.trace_mode == 0 && tid == 0) tc_trace(p, tick, 0);
        const int t = tick - sigma;
        const bool act = warp_on && s < B && t >= 0 && t < len;
        const float4 *base = logits4 + (size_t)((tick + 1 - p.zlag) & p.zmask) * Vq * SB + s;
        float tl = 0.f;
        if (act && sl == 0 && !full_rows) {
            int tg = __ldg(targets + o0 + t);
            tg = min(max(tg, 0), V - 1);
            tl = __ldcg(reinterpret_cast<const float *>(base + (size_t)(tg >> 2) * SB) + (tg & 3));
        }
        float m = -INFINITY, sum = 0.f;
        if (warp_on && t >= 0) {
            for (int v0 = v4_lo; v0 < v4_hi; v0 += 8) {
                float4 x[8];
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    x[j] = (v0 + j < v4_hi) ? __ldcg(base + (size_t)(v0 + j) * SB) : make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
                if (4 * (v0 + 8) > V) {                      // mask the padding entries of the last float4
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int v = 4 * (v0 + j);
                        if (v + 1 >= V) x[j].y = -INFINITY;
                        if (v + 2 >= V) x[j].z = -INFINITY;
                        if (v + 3 >= V) x[j].w = -INFINITY;
                        if (v >= V) x[j].x = -INFINITY;
                    }
                }
                float mx = m;
#pragma unroll
                for (int j = 0; j < 8; ++j) mx = fmaxf(fmaxf(fmaxf(mx, x[j].x), fmaxf(x[j].y, x[j].z)), x[j].w);
                if (mx > -INFINITY) {
                    float acc = sum * __expf(m - mx);
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        acc += __expf(x[j].x - mx) + __expf(x[j].y - mx) + __expf(x[j].z - mx) + __expf(x[j].w - mx);
                    sum = acc; m = mx;
                }
            }
        }
        part_m[warp][lane] = m; part_s[warp][lane] = sum;
        __syncthreads();
        // every warp of the group combines the slices (cheap), slice 0 writes the score
        float M = -INFINITY, S = 0.f;
        if (warp_on) {
            for (int i = 0; i < nslices; ++i) M = fmaxf(M, part_m[gl * nslices + i][lane]);
            for (int i = 0; i < nslices; ++i) {
                const float mi = part_m[gl * nslices + i][lane];
                if (mi > -INFINITY) S += part_s[gl * nslices + i][lane] * __expf(mi - M);
            }
        }
        if (act) {
            if (!full_rows) {
                if (sl == 0) out[o0 + t] = (tl - M) - logf(S);
            } else {
                float *dst = out + (size_t)(o0 + t) * V;
                const float inv = 1.0f / S;
                for (int v4 = v4_lo; v4 < v4_hi; ++v4) {
                    const float4 x = __ldcg(base + (size_t)v4 * SB);
                    const int v = 4 * v4;
                    dst[v] = expf(x.x - M) * inv;
                    if (v + 1 < V) dst[v + 1] = expf(x.y - M) * inv;
                    if (v + 2 < V) dst[v + 2] = expf(x.z - M) * inv;
                    if (v + 3 < V) dst[v + 3] = expf(x.w - M) * inv;
                }
            }
        }
        __syncthreads();
        if (tid == 0) {
            tc::red_release_add(p.barrier, 1u);
            if (p.trace && p.trace_mode == 0) tc_trace(p, tick, 7);
        }
    }
}
