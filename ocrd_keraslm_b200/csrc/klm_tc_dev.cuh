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
static __device__ void tc_lse_role(const KlmTcParams &p, const KlmTcRole &role) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const unsigned G = gridDim.x;
    const int nticks = p.T_max + 2 * p.D;
    const int nwarps = KLM_TC_THREADS / 32;
    __shared__ int ok_s;
    if (tid == 0) tc::red_release_add(p.barrier, 1u);
    for (int tick = 0; tick < nticks; ++tick) {
        if (tid == 0) ok_s = tc_wait_grid(p, (unsigned)(tick + 1) * G, 10) ? 1 : 0;
        __syncthreads();
        if (!ok_s) return;
        const int t = tick - role.sigma;
        if (t >= 0) {
            for (int s = role.first_pair * nwarps + warp; s < p.B; s += role.npairs * nwarps) {
                const long long o0 = p.offsets[s];
                const int len = (int)(p.offsets[s + 1] - o0);
                if (t >= len) continue;
                const float *row = p.logits + ((size_t)(tick & 1) * 64 + s) * p.Vp;
                // one pass: running maximum and rescaled sum per lane, then combined across the warp
                float m = -INFINITY, sum = 0.f;
                for (int v0 = 0; v0 < p.V; v0 += 256) {
                    float x[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int v = v0 + 32 * j + lane;
                        x[j] = v < p.V ? __ldcg(row + v) : -INFINITY;
                    }
                    float mx = m;
#pragma unroll
                    for (int j = 0; j < 8; ++j) mx = fmaxf(mx, x[j]);
                    if (mx > -INFINITY) {
                        float acc = sum * __expf(m - mx);
#pragma unroll
                        for (int j = 0; j < 8; ++j) acc += __expf(x[j] - mx);
                        sum = acc; m = mx;
                    }
                }
                float mw = m;
                for (int o = 16; o > 0; o >>= 1) mw = fmaxf(mw, __shfl_xor_sync(0xffffffffu, mw, o));
                sum = (m > -INFINITY) ? sum * __expf(m - mw) : 0.f;
                for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
                m = mw;
                if (p.full_rows) {
                    float *dst = p.out + (size_t)(o0 + t) * p.V;
                    const float inv = 1.0f / sum;
                    for (int v = lane; v < p.V; v += 32) dst[v] = expf(__ldcg(row + v) - m) * inv;
                } else if (lane == 0) {
                    int tg = __ldg(p.targets + o0 + t);
                    tg = min(max(tg, 0), p.V - 1);
                    p.out[o0 + t] = (__ldcg(row + tg) - m) - logf(sum);
                }
            }
        }
        __syncthreads();
        if (tid == 0) tc::red_release_add(p.barrier, 1u);
    }
}

