// Device helpers shared by the tensor-core sequence kernels (klm_seq_tc.cu, klm_seq_tc2.cu):
// bounded waits with a device-side watchdog, timeline tracing, the LSTM cell, the LSE stage.
#pragma once
#include "klm_common.cuh"
#include "klm_tc.cuh"

#define TC_SPIN_LIMIT (1u << 24)

__device__ __forceinline__ bool tc_spin_fail(unsigned &n, const KlmTcParams &p, int code) {
    if ((++n & 0x3FFu) == 0) {
        if (*(volatile int *)p.err != 0) return true;
        if (n > TC_SPIN_LIMIT) { atomicCAS(p.err, 0, code); return true; }
    }
    return false;
}
__device__ __forceinline__ bool tc_wait_grid(const KlmTcParams &p, unsigned target, int code) {
    unsigned n = 0;
    while (tc::ld_acquire(p.barrier) < target)
        if (tc_spin_fail(n, p, code)) return false;
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
        if (k >= 0 && k < KLM_TC_TRACE_TICKS)
            p.trace[((size_t)blockIdx.x * KLM_TC_TRACE_TICKS + k) * 8 + ev] = (unsigned long long)clock64();
    }
}
__device__ __forceinline__ void tc_bar_epi() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// exp via ex2.approx (2 ulp), reciprocal via rcp.approx: absolute error of the gates ~1e-7
__device__ __forceinline__ float tc_exp(float x) { return exp2f(x * 1.4426950408889634f); }
__device__ __forceinline__ float tc_clamp(float x) { return fminf(fmaxf(x, -30.f), 30.f); }

// LSTM cell (Keras LSTMCell, gate order i|f|c|o, exact logistic; rating.py:133-134), 8 MUFU ops:
//   i*g = (1 - e^-2g) / ((1 + e^-i)(1 + e^-2g)),  f = 1/(1 + e^-f),  h = o*tanh(c) likewise
__device__ __forceinline__ void tc_cell(float ai, float af, float ag, float ao, float &c, float &h) {
    const float ei = tc_exp(-tc_clamp(ai)), ef = tc_exp(-tc_clamp(af)), eg = tc_exp(-2.f * tc_clamp(ag)),
                eo = tc_exp(-tc_clamp(ao));
    const float ig = __fdividef(1.f - eg, (1.f + ei) * (1.f + eg));
    const float f = __fdividef(1.f, 1.f + ef);
    c = fmaf(f, c, ig);
    const float ec = tc_exp(-2.f * tc_clamp(c));
    h = __fdividef(1.f - ec, (1.f + eo) * (1.f + ec));
}

__device__ __forceinline__ void tc_split(float v, __half &hi, __half &lo) {
    hi = __float2half_rn(v);
    lo = __float2half_rn(v - __half2float(hi));
}

// ---------------------------------------------------------------------------------------------
// LSE stage: log-sum-exp over the logits of every stream + pick of the target's logit (or the full softmax row).
// A warp owns up to 32 streams (lane k keeps stream k's offsets); per tick it walks over (stream, 256-logit block)
// pairs with the loads of the next block in flight while the current one is reduced, and the target index and
// target logit of all its streams are fetched up front by one lane each -- per tick the stage costs about one L2
// round trip plus ~150 clk per stream instead of three dependent round trips per stream.
static __device__ void tc_lse_role(const KlmTcParams &p, const KlmTcRole &role) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const unsigned G = gridDim.x;
    const int nticks = p.T_max + 2 * p.D;
    const int nwarps = KLM_TC_THREADS / 32;
    const int V = p.V, Vp = p.Vp, B = p.B, sigma = role.sigma, full_rows = p.full_rows;
    const float *logits = p.logits;
    const int *targets = p.targets;
    float *out = p.out;
    __shared__ int ok_s;
    // my streams: first, first + stride, ...; lane k holds stream k's data
    const int first = role.first_pair * nwarps + warp, stride = role.npairs * nwarps;
    const int nmine = first < B ? (B - first + stride - 1) / stride : 0;           // <= 32 (B <= 64, stride >= 10)
    const int s_lane = first + lane * stride;
    long long o0_lane = 0;
    int len_lane = 0;
    if (lane < nmine) { o0_lane = p.offsets[s_lane]; len_lane = (int)(p.offsets[s_lane + 1] - o0_lane); }
    const int nblk = (V + 255) / 256, total = nmine * nblk;
    if (tid == 0) tc::red_release_add(p.barrier, 1u);
    for (int tick = 0; tick < nticks; ++tick) {
        if (tid == 0) ok_s = tc_wait_grid(p, (unsigned)(tick + 1) * G, 10) ? 1 : 0;
        __syncthreads();
        if (!ok_s) return;
        if (p.trace && p.trace_mode == 0 && tid == 0) tc_trace(p, tick, 0);
        const int t = tick - sigma;
        if (t >= 0 && total > 0) {
            const float *base = logits + (size_t)(tick & 1) * 64 * Vp;
            // target logit of my streams, one lane per stream
            const bool act_lane = lane < nmine && t < len_lane;
            float tl_lane = 0.f;
            if (act_lane && !full_rows) {
                int tg = __ldg(targets + o0_lane + t);
                tg = min(max(tg, 0), V - 1);
                tl_lane = __ldcg(base + (size_t)s_lane * Vp + tg);
            }
            if (nblk == 1 && nmine <= 8) {
                // common case (V <= 256, at most 8 streams per warp): all loads in flight at once
                float x[8][8];
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    if (k < nmine) {
                        const float *row = base + (size_t)(first + k * stride) * Vp;
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const int v = 32 * j + lane;
                            x[k][j] = v < V ? __ldcg(row + v) : -INFINITY;
                        }
                    }
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    if (k < nmine) {
                        float m = x[k][0];
#pragma unroll
                        for (int j = 1; j < 8; ++j) m = fmaxf(m, x[k][j]);
                        for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
                        float sum = 0.f;
#pragma unroll
                        for (int j = 0; j < 8; ++j) sum += __expf(x[k][j] - m);
                        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
                        const int act = __shfl_sync(0xffffffffu, act_lane ? 1 : 0, k);
                        if (full_rows) {
                            const long long o0 = __shfl_sync(0xffffffffu, o0_lane, k);
                            if (act) {
                                float *dst = out + (size_t)(o0 + t) * V;
                                const float inv = 1.0f / sum;
#pragma unroll
                                for (int j = 0; j < 8; ++j) {
                                    const int v = 32 * j + lane;
                                    if (v < V) dst[v] = expf(x[k][j] - m) * inv;
                                }
                            }
                        } else if (lane == k && act) {
                            out[o0_lane + t] = (tl_lane - m) - logf(sum);
                        }
                    }
                goto lse_tick_done;
            }
            float cur[8], nxt[8];
            auto load_blk = [&](int b, float (&x)[8]) {
                const int k = b / nblk, v0 = (b - k * nblk) * 256;
                const float *row = base + (size_t)(first + k * stride) * Vp;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int v = v0 + 32 * j + lane;
                    x[j] = v < V ? __ldcg(row + v) : -INFINITY;
                }
            };
            load_blk(0, cur);
            float m = -INFINITY, sum = 0.f;
            for (int b = 0; b < total; ++b) {
                if (b + 1 < total) load_blk(b + 1, nxt);
                const int k = b / nblk, blk = b - k * nblk;
                float mx = m;
#pragma unroll
                for (int j = 0; j < 8; ++j) mx = fmaxf(mx, cur[j]);
                if (mx > -INFINITY) {
                    float acc = sum * __expf(m - mx);
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc += __expf(cur[j] - mx);
                    sum = acc; m = mx;
                }
                if (blk == nblk - 1) {                      // stream k complete: combine across the warp
                    float mw = m;
                    for (int o = 16; o > 0; o >>= 1) mw = fmaxf(mw, __shfl_xor_sync(0xffffffffu, mw, o));
                    sum = (m > -INFINITY) ? sum * __expf(m - mw) : 0.f;
                    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
                    const int act = __shfl_sync(0xffffffffu, act_lane ? 1 : 0, k);
                    if (full_rows) {
                        const long long o0 = __shfl_sync(0xffffffffu, o0_lane, k);
                        if (act) {
                            const float *row = base + (size_t)(first + k * stride) * Vp;
                            float *dst = out + (size_t)(o0 + t) * V;
                            const float inv = 1.0f / sum;
                            for (int v = lane; v < V; v += 32) dst[v] = expf(__ldcg(row + v) - mw) * inv;
                        }
                    } else if (lane == k && act) {
                        out[o0_lane + t] = (tl_lane - mw) - logf(sum);
                    }
                    m = -INFINITY; sum = 0.f;
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) cur[j] = nxt[j];
            }
        }
    lse_tick_done:
        __syncthreads();
        if (tid == 0) {
            tc::red_release_add(p.barrier, 1u);
            if (p.trace && p.trace_mode == 0) tc_trace(p, tick, 7);
        }
    }
}

