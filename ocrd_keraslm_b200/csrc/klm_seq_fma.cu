// Persistent stacked-LSTM sequence kernel, FP32 FMA path (small numbers of streams / hypotheses,
// and the general fallback shape for any model that the tensor-core path does not take).
//
// ONE cooperative launch streams whole character sequences: fused embedding gather (as a lookup
// into the pre-contracted tables P/Q), depth x LSTM, tied output projection, log-softmax and
// target pick.  Replaces the per-window Session.run of Rater.rate (rating.py:516), the
// per-character one of Rater.rate2 (:566) and the per-step one of Rater.predict (:631).
//
// Decomposition: every CTA owns a fixed slice of one stage for the whole launch --
//   LSTM layer l : `count` units (= 4 gate columns each), weights resident in shared memory
//   PROJ         : `count` groups of 4 vocabulary rows of E
//   LSE          : one CTA, log-sum-exp over the logits and the target pick / full softmax rows
// and the stages run as a wavefront over "ticks": at tick k layer l works on position k-l,
// PROJ on k-D, LSE on k-D-1.  Hidden vectors travel through L2 in hbuf[l][parity][unit][stream];
// one grid-wide barrier per tick.  Cell states never leave the registers of their owner thread.
#include <cstdio>
#include <cstring>
#include "klm_common.cuh"

__device__ __forceinline__ float klm_sigmoid(float x) { return 1.0f / (1.0f + expf(-x)); }

__device__ __forceinline__ unsigned int ld_acquire_u32(const unsigned int *p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Grid-wide barrier on a monotonically increasing counter (kernel is launched cooperatively,
// so all CTAs are co-resident).
__device__ __forceinline__ void grid_barrier(unsigned int *ctr, unsigned int target) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(ctr, 1u);
        while (ld_acquire_u32(ctr) < target) { }
        __threadfence();
    }
    __syncthreads();
}

__device__ __forceinline__ void fma4(float4 &a, const float4 w, const float h) {
    a.x = fmaf(w.x, h, a.x); a.y = fmaf(w.y, h, a.y); a.z = fmaf(w.z, h, a.z); a.w = fmaf(w.w, h, a.w);
}
__device__ __forceinline__ void add4(float4 &a, const float4 b) { a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }

// Pre-activation of layer-1 inputs: gather of the pre-contracted embedding rows
// (fused char + context embedding gather/concat + input contraction; rating.py:103-115,125).
__device__ __forceinline__ float4 gather_input(const KlmSeqParams &p, int b, long long pos, int t, int u) {
    int x = __ldg(p.chars + pos);
    x = min(max(x, 0), p.V - 1);
    float4 pre = __ldg(p.P + (size_t)x * p.W + u);
    const bool real = p.n_real ? (t < __ldg(p.n_real + b)) : true;
    for (int k = 0; k < p.Nc; ++k) {
        int z = real ? __ldg(p.ctx + b * p.Nc + k) : 0;       // padding: context 0 (:1141-1143)
        z = min(max(z, 0), p.ctx_rows[k] - 1);
        add4(pre, __ldg(p.Q[k] + (size_t)z * p.W + u));
    }
    return pre;
}

// LSTM cell update (Keras LSTMCell, gate order i|f|c|o, exact sigmoid; rating.py:133-134).
__device__ __forceinline__ void lstm_cell(const float4 a, float &c, float &h) {
    const float i = klm_sigmoid(a.x), f = klm_sigmoid(a.y), g = tanhf(a.z), o = klm_sigmoid(a.w);
    c = fmaf(f, c, i * g);
    h = o * tanhf(c);
}

// Same cell with 8 MUFU operations (ex2.approx / rcp.approx, arguments clamped to +-30; absolute error of the gates
// ~1e-7, the formulation of the tensor-core kernel): a single lane runs it on the recurrence of the barrier-free
// kernel, where the ~100 dependent instructions of expf/tanhf/division are pure latency.
__device__ __forceinline__ float klm_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x * 1.4426950408889634f));
    return y;
}
__device__ __forceinline__ float klm_clamp30(float x) { return fminf(fmaxf(x, -30.f), 30.f); }
__device__ __forceinline__ void lstm_cell_fast(const float4 a, float &c, float &h) {
    const float ei = klm_ex2(-klm_clamp30(a.x)), ef = klm_ex2(-klm_clamp30(a.y)), eg = klm_ex2(-2.f * klm_clamp30(a.z)),
                eo = klm_ex2(-klm_clamp30(a.w));
    float num, den;
    klm_tanh_parts(a.z, eg, num, den);
    const float ig = __fdividef(num, (1.f + ei) * den);
    const float f = __fdividef(1.f, 1.f + ef);
    c = fmaf(f, c, ig);
    const float ec = klm_ex2(-2.f * klm_clamp30(c));
    klm_tanh_parts(c, ec, num, den);
    h = __fdividef(num, (1.f + eo) * den);
}

// ---------------------------------------------------------------------------------------------
// LSE stage: one warp per stream, lanes over the vocabulary.
__device__ __forceinline__ void lse_stage(const KlmSeqParams &p, int tick) {
    const int t = tick - p.D - 1;
    if (t < 0) return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float *lg = p.logits + (size_t)(tick & 1) * p.V4 * p.Bp;
    for (int b = warp; b < p.B; b += KLM_WARPS) {
        const long long o0 = p.offsets[b];
        const int len = (int)(p.offsets[b + 1] - o0);
        if (t >= len) continue;
        float m = -INFINITY;
        for (int v = lane; v < p.V; v += 32) m = fmaxf(m, __ldcg(lg + (size_t)v * p.Bp + b));
        for (int s = 16; s > 0; s >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, s));
        float sum = 0.f;
        for (int v = lane; v < p.V; v += 32) sum += expf(__ldcg(lg + (size_t)v * p.Bp + b) - m);
        for (int s = 16; s > 0; s >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, s);
        if (p.full_rows) {
            float *row = p.out + (size_t)(o0 + t) * p.V;
            const float inv = 1.0f / sum;
            for (int v = lane; v < p.V; v += 32) row[v] = expf(__ldcg(lg + (size_t)v * p.Bp + b) - m) * inv;
        } else if (lane == 0) {
            int tg = __ldg(p.targets + o0 + t);
            tg = min(max(tg, 0), p.V - 1);
            p.out[o0 + t] = (__ldcg(lg + (size_t)tg * p.Bp + b) - m) - logf(sum);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Stream-parallel mode (B > 8): lane = stream, each thread walks K sequentially; the hidden vectors
// are staged [k][stream] so that the weight read is a broadcast and the h read is conflict-free.
__global__ void __launch_bounds__(KLM_THREADS, 1) klm_seq_fma_streams(const KlmSeqParams p) {
    extern __shared__ float4 smem4[];
    const KlmRole role = p.roles[blockIdx.x];
    const int W = p.W, B = p.B, Bp = p.Bp;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int l = role.layer;
    const bool is_lstm = role.kind == KLM_ROLE_LSTM, is_proj = role.kind == KLM_ROLE_PROJ;
    const int K = (is_lstm && l > 0) ? 2 * W : W;
    const int nsg = Bp / 32;
    const int nitems = role.count * nsg;

    const float4 *wptr = nullptr;
    float *stage = reinterpret_cast<float *>(smem4);
    if (is_lstm || is_proj) {
        wptr = (is_lstm ? p.Wl[l] : p.Ep) + (size_t)role.first * K;
        if (p.wsmem) {
            const int n4 = role.count * K;
            for (int i = threadIdx.x; i < n4; i += KLM_THREADS) smem4[i] = __ldg(wptr + i);
            wptr = smem4;
            stage = reinterpret_cast<float *>(smem4 + n4);
        }
    }
    const int KC = p.stage_floats / Bp;

    float c_reg[KLM_MAX_ITEMS], h_reg[KLM_MAX_ITEMS];
    // ---- init: load the carried state of own units, publish h into parity 0 ----
#pragma unroll
    for (int j = 0; j < KLM_MAX_ITEMS; ++j) {
        c_reg[j] = 0.f; h_reg[j] = 0.f;
        const int it = warp + j * KLM_WARPS;
        if (is_lstm && it < nitems) {
            const int u = role.first + it / nsg, b = (it % nsg) * 32 + lane;
            if (b < B) {
                const int slot = p.slots_in[b];
                if (slot >= 0) {
                    const float *st = p.state + ((size_t)slot * p.D + l) * 2 * W;
                    h_reg[j] = st[u]; c_reg[j] = st[W + u];
                }
                p.hbuf[((size_t)(l * 2 + 0) * W + u) * Bp + b] = h_reg[j];
            }
        }
    }
    unsigned int bar = 0;
    grid_barrier(p.barrier, (++bar) * gridDim.x);

    const int nticks = p.T_max + p.D + 1;
    for (int tick = 0; tick < nticks; ++tick) {
        const int par = tick & 1;
        if (is_lstm || is_proj) {
            float4 acc[KLM_MAX_ITEMS];
#pragma unroll
            for (int j = 0; j < KLM_MAX_ITEMS; ++j) acc[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            const int nsrc = (is_lstm && l > 0) ? 2 : 1;
            for (int src = 0; src < nsrc; ++src) {
                int hl;                                      // which layer's hidden vector
                if (is_proj) hl = p.D - 1; else if (l == 0) hl = 0; else hl = (src == 0) ? l - 1 : l;
                const float *hsrc = p.hbuf + (size_t)(hl * 2 + par) * W * Bp;
                for (int k0 = 0; k0 < W; k0 += KC) {
                    const int kc = min(KC, W - k0);
                    __syncthreads();
                    const float4 *g4 = reinterpret_cast<const float4 *>(hsrc + (size_t)k0 * Bp);
                    float4 *s4 = reinterpret_cast<float4 *>(stage);
                    for (int i = threadIdx.x; i < kc * Bp / 4; i += KLM_THREADS) s4[i] = __ldcg(g4 + i);
                    __syncthreads();
#pragma unroll
                    for (int j = 0; j < KLM_MAX_ITEMS; ++j) {
                        const int it = warp + j * KLM_WARPS;
                        if (it < nitems) {
                            const float4 *wrow = wptr + (size_t)(it / nsg) * K + src * W + k0;
                            const float *hcol = stage + (it % nsg) * 32 + lane;
                            float4 a = acc[j];
#pragma unroll 8
                            for (int k = 0; k < kc; ++k) fma4(a, wrow[k], hcol[k * Bp]);
                            acc[j] = a;
                        }
                    }
                }
            }
            // ---- epilogue ----
#pragma unroll
            for (int j = 0; j < KLM_MAX_ITEMS; ++j) {
                const int it = warp + j * KLM_WARPS;
                if (it < nitems) {
                    const int ui = role.first + it / nsg, b = (it % nsg) * 32 + lane;
                    if (b < B) {
                        if (is_lstm) {
                            const int t = tick - l;
                            const long long o0 = p.offsets[b];
                            const int len = (int)(p.offsets[b + 1] - o0);
                            if (t >= 0 && t < len) {
                                float4 a = acc[j];
                                if (l == 0) add4(a, gather_input(p, b, o0 + t, t, ui));
                                else add4(a, __ldg(p.bias[l] + ui));
                                lstm_cell(a, c_reg[j], h_reg[j]);
                            }
                            p.hbuf[((size_t)(l * 2 + (par ^ 1)) * W + ui) * Bp + b] = h_reg[j];
                        } else {
                            float *lg = p.logits + ((size_t)(par ^ 1) * p.V4 + 4 * ui) * Bp + b;
                            lg[0] = acc[j].x; lg[Bp] = acc[j].y; lg[2 * Bp] = acc[j].z; lg[3 * Bp] = acc[j].w;
                        }
                    }
                }
            }
        } else if (role.kind == KLM_ROLE_LSE) {
            lse_stage(p, tick);
        }
        grid_barrier(p.barrier, (++bar) * gridDim.x);
    }

    // ---- write the final state back ----
    if (is_lstm) {
#pragma unroll
        for (int j = 0; j < KLM_MAX_ITEMS; ++j) {
            const int it = warp + j * KLM_WARPS;
            if (it < nitems) {
                const int u = role.first + it / nsg, b = (it % nsg) * 32 + lane;
                if (b < B) {
                    float *st = p.state + ((size_t)p.slots_out[b] * p.D + l) * 2 * W;
                    st[u] = h_reg[j]; st[W + u] = c_reg[j];
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// K-parallel mode (B <= 8: single documents, tiny beams): lanes split the contraction index,
// partial sums are combined with warp shuffles; the whole input vector of a tick is staged once.
template <int BT>
__global__ void __launch_bounds__(KLM_THREADS, 1) klm_seq_fma_kpar(const KlmSeqParams p) {
    extern __shared__ float4 smem4[];
    const KlmRole role = p.roles[blockIdx.x];
    const int W = p.W, B = p.B, Bp = p.Bp;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int l = role.layer;
    const bool is_lstm = role.kind == KLM_ROLE_LSTM, is_proj = role.kind == KLM_ROLE_PROJ;
    const int K = (is_lstm && l > 0) ? 2 * W : W;

    const float4 *wptr = nullptr;
    float *stage = reinterpret_cast<float *>(smem4);
    if (is_lstm || is_proj) {
        wptr = (is_lstm ? p.Wl[l] : p.Ep) + (size_t)role.first * K;
        if (p.wsmem) {
            const int n4 = role.count * K;
            for (int i = threadIdx.x; i < n4; i += KLM_THREADS) smem4[i] = __ldg(wptr + i);
            wptr = smem4;
            stage = reinterpret_cast<float *>(smem4 + n4);
        }
    }

    // lane b (< B) of the warp that owns unit j keeps that stream's cell/hidden value
    float c_reg[KLM_MAX_UNITS_K], h_reg[KLM_MAX_UNITS_K];
#pragma unroll
    for (int j = 0; j < KLM_MAX_UNITS_K; ++j) {
        c_reg[j] = 0.f; h_reg[j] = 0.f;
        const int ui = warp + j * KLM_WARPS;
        if (is_lstm && ui < role.count && lane < B) {
            const int u = role.first + ui, b = lane;
            const int slot = p.slots_in[b];
            if (slot >= 0) {
                const float *st = p.state + ((size_t)slot * p.D + l) * 2 * W;
                h_reg[j] = st[u]; c_reg[j] = st[W + u];
            }
            p.hbuf[((size_t)(l * 2 + 0) * W + u) * Bp + b] = h_reg[j];
        }
    }
    unsigned int bar = 0;
    grid_barrier(p.barrier, (++bar) * gridDim.x);

    const int nticks = p.T_max + p.D + 1;
    for (int tick = 0; tick < nticks; ++tick) {
        const int par = tick & 1;
        if (is_lstm || is_proj) {
            // ---- stage the input vector(s): stage[b][K] <- hbuf[hl][par][k][b] ----
            const int nsrc = (is_lstm && l > 0) ? 2 : 1;
            for (int src = 0; src < nsrc; ++src) {
                int hl;
                if (is_proj) hl = p.D - 1; else if (l == 0) hl = 0; else hl = (src == 0) ? l - 1 : l;
                const float *hsrc = p.hbuf + (size_t)(hl * 2 + par) * W * Bp;
                for (int i = threadIdx.x; i < W * B; i += KLM_THREADS) {
                    const int k = i / B, b = i - k * B;
                    stage[b * K + src * W + k] = __ldcg(hsrc + (size_t)k * Bp + b);
                }
            }
            __syncthreads();
#pragma unroll
            for (int j = 0; j < KLM_MAX_UNITS_K; ++j) {
                const int ui = warp + j * KLM_WARPS;
                if (ui < role.count) {
                    float4 acc[BT];
#pragma unroll
                    for (int bb = 0; bb < BT; ++bb) acc[bb] = make_float4(0.f, 0.f, 0.f, 0.f);
                    const float4 *wrow = wptr + (size_t)ui * K;
                    for (int k = lane; k < K; k += 32) {
                        const float4 w = wrow[k];
#pragma unroll
                        for (int bb = 0; bb < BT; ++bb)
                            if (bb < B) fma4(acc[bb], w, stage[bb * K + k]);
                    }
                    float4 mine = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                    for (int bb = 0; bb < BT; ++bb) {
                        // 4 sums over 32 lanes in 6 shuffles: halve the set of values a lane carries while folding
                        const bool up16 = lane & 16, up8 = lane & 8;
                        float s0 = up16 ? acc[bb].x : acc[bb].z, s1 = up16 ? acc[bb].y : acc[bb].w;   // what the partner keeps
                        float k0 = up16 ? acc[bb].z : acc[bb].x, k1 = up16 ? acc[bb].w : acc[bb].y;   // what this lane keeps
                        k0 += __shfl_xor_sync(0xffffffffu, s0, 16);
                        k1 += __shfl_xor_sync(0xffffffffu, s1, 16);
                        float v = (up8 ? k1 : k0) + __shfl_xor_sync(0xffffffffu, up8 ? k0 : k1, 8);
                        v += __shfl_xor_sync(0xffffffffu, v, 4);
                        v += __shfl_xor_sync(0xffffffffu, v, 2);
                        v += __shfl_xor_sync(0xffffffffu, v, 1);
                        // lanes 0-7 hold x, 8-15 y, 16-23 z, 24-31 w: collect them in lane bb
                        const float gx = __shfl_sync(0xffffffffu, v, 0), gy = __shfl_sync(0xffffffffu, v, 8),
                                    gz = __shfl_sync(0xffffffffu, v, 16), gw = __shfl_sync(0xffffffffu, v, 24);
                        if (lane == bb) mine = make_float4(gx, gy, gz, gw);
                    }
                    if (lane < B) {
                        const int u = role.first + ui, b = lane;
                        if (is_lstm) {
                            const int t = tick - l;
                            const long long o0 = p.offsets[b];
                            const int len = (int)(p.offsets[b + 1] - o0);
                            if (t >= 0 && t < len) {
                                if (l == 0) add4(mine, gather_input(p, b, o0 + t, t, u));
                                else add4(mine, __ldg(p.bias[l] + u));
                                lstm_cell(mine, c_reg[j], h_reg[j]);
                            }
                            p.hbuf[((size_t)(l * 2 + (par ^ 1)) * W + u) * Bp + b] = h_reg[j];
                        } else {
                            float *lg = p.logits + ((size_t)(par ^ 1) * p.V4 + 4 * u) * Bp + b;
                            lg[0] = mine.x; lg[Bp] = mine.y; lg[2 * Bp] = mine.z; lg[3 * Bp] = mine.w;
                        }
                    }
                }
            }
        } else if (role.kind == KLM_ROLE_LSE) {
            lse_stage(p, tick);
        }
        grid_barrier(p.barrier, (++bar) * gridDim.x);
    }

    if (is_lstm) {
#pragma unroll
        for (int j = 0; j < KLM_MAX_UNITS_K; ++j) {
            const int ui = warp + j * KLM_WARPS;
            if (ui < role.count && lane < B) {
                const int u = role.first + ui;
                float *st = p.state + ((size_t)p.slots_out[lane] * p.D + l) * 2 * W;
                st[u] = h_reg[j]; st[W + u] = c_reg[j];
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// K-parallel mode without a grid barrier (B <= 8; the default for single documents and tiny beams).
//
// With one stream the tick is pure latency, and the barrier version pays three dependent L2 round trips per
// tick: hidden-vector stores + fence, the counter (atomic + poll), then the loads of the next tick.  Here every
// exchanged value travels as ONE 8-byte word {value, tick it is meant for} (the "LL" scheme of collective
// libraries: an aligned 8-byte store is a single transaction, so a reader that sees the tick number has the
// value): consumers simply poll the words they need, there is no fence, no counter and no barrier on the path.
// The words live in a ring of KLM_LL_RING ticks; every KLM_LL_RING/2 ticks a CTA checks (one coalesced read of
// per-CTA progress words, almost never a wait) that nobody is more than half a ring behind before slots are reused.
// The ring is filled with 0xFF by the host before each launch, so stale words never match.
#ifndef KLM_LL_WEAK_STORE
#define KLM_LL_WEAK_STORE 1
#endif
__device__ __forceinline__ uint2 ld_volatile_u2(const uint2 *p) {
    uint2 v;
    asm volatile("ld.relaxed.gpu.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_u2(uint2 *p, unsigned x, unsigned y) {
    if (KLM_LL_WEAK_STORE) asm volatile("st.global.cg.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(x), "r"(y) : "memory");
    else asm volatile("st.relaxed.gpu.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(x), "r"(y) : "memory");
}
// value of a word once it carries `tick`; gives up (and flags the launch) after ~2^22 polls
__device__ __forceinline__ float ll_wait(const uint2 *p, unsigned tick, int *err) {
    uint2 v = ld_volatile_u2(p);
    unsigned n = 0;
    while (v.y != tick) {
        if (++n > (1u << 22)) { atomicCAS(err, 0, 20); break; }
        if ((n & 0xFFu) == 0 && *(volatile int *)err != 0) break;
        v = ld_volatile_u2(p);
    }
    return __uint_as_float(v.x);
}

template <int BT>
__global__ void __launch_bounds__(KLM_THREADS, 1) klm_seq_fma_ll(const __grid_constant__ KlmSeqParams p) {
    extern __shared__ float4 smem4[];
    const KlmRole role = p.roles[blockIdx.x];
    const int W = p.W, B = p.B, D = p.D;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int l = role.layer;
    const bool is_lstm = role.kind == KLM_ROLE_LSTM, is_proj = role.kind == KLM_ROLE_PROJ;
    const int K = (is_lstm && l > 0) ? 2 * W : W;
    constexpr int R = KLM_LL_RING;                           // ring slots
    const int BS = B;                                        // words of one unit: one per stream, contiguous (a warp poll = 256 bytes)
    uint2 *hx = p.hbuf_ll;                                   // [R][D][W][B]
    uint2 *lgx = p.logits_ll;                                // [R][4*V4][B]

    const float4 *wptr = nullptr;
    float *stage = reinterpret_cast<float *>(smem4);
    if (is_lstm || is_proj) {
        wptr = (is_lstm ? p.Wl[l] : p.Ep) + (size_t)role.first * K;
        if (p.wsmem) {
            const int n4 = role.count * K;
            for (int i = threadIdx.x; i < n4; i += KLM_THREADS) smem4[i] = __ldg(wptr + i);
            wptr = smem4;
            stage = reinterpret_cast<float *>(smem4 + n4);
        }
    }

    __shared__ float outbuf[KLM_WARPS * KLM_MAX_UNITS_K * 4 * KLM_MODEK_MAXB];      // this tick's results before they leave
    float c_reg[KLM_MAX_UNITS_K], h_reg[KLM_MAX_UNITS_K];
    long long o0 = 0;
    int len = 0;
    if (lane < B) { o0 = p.offsets[lane]; len = (int)(p.offsets[lane + 1] - o0); }
#pragma unroll
    for (int j = 0; j < KLM_MAX_UNITS_K; ++j) {
        c_reg[j] = 0.f; h_reg[j] = 0.f;
        const int ui = warp + j * KLM_WARPS;
        if (is_lstm && ui < role.count && lane < B) {
            const int u = role.first + ui, b = lane;
            const int slot = p.slots_in[b];
            if (slot >= 0) {
                const float *st = p.state + ((size_t)slot * D + l) * 2 * W;
                h_reg[j] = st[u]; c_reg[j] = st[W + u];
            }
            st_volatile_u2(hx + ((size_t)(0 * D + l) * W + u) * BS + b, __float_as_uint(h_reg[j]), 0u);
        }
    }

    int tg_next = 0, tg_next_t = -1;                         // LSE: target index of the next position (stream = warp)
    // layer 1: nothing of the embedding gather may sit on the recurrence -- the character index is fetched a tick
    // ahead, the context part sum_k Q_k[z_k][u] is tick-invariant, and the table rows are read BEFORE polling for h
    int x_next = 0, nreal = 0x7fffffff;
    float4 qsum[KLM_MAX_UNITS_K], pre[KLM_MAX_UNITS_K];
#pragma unroll
    for (int j = 0; j < KLM_MAX_UNITS_K; ++j) qsum[j] = pre[j] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (is_lstm && lane < B) {
        if (l == 0) {
            if (len > 0) x_next = __ldg(p.chars + o0);
            if (p.n_real) nreal = __ldg(p.n_real + lane);
            for (int k = 0; k < p.Nc; ++k) {
                int z = __ldg(p.ctx + lane * p.Nc + k);
                z = min(max(z, 0), p.ctx_rows[k] - 1);
#pragma unroll
                for (int j = 0; j < KLM_MAX_UNITS_K; ++j) {
                    const int ui = warp + j * KLM_WARPS;
                    if (ui < role.count) add4(qsum[j], __ldg(p.Q[k] + (size_t)z * W + role.first + ui));
                }
            }
        } else {
#pragma unroll
            for (int j = 0; j < KLM_MAX_UNITS_K; ++j) {
                const int ui = warp + j * KLM_WARPS;
                if (ui < role.count) pre[j] = __ldg(p.bias[l] + role.first + ui);
            }
        }
    }
    const int nticks = p.T_max + D + 1;
    long long dbg_wait = 0, dbg_work = 0, dbg_t0 = 0, dbg_ph[6] = {0, 0, 0, 0, 0, 0}, dbg_t1 = 0;
#define LL_DBG(i) if (p.debug > 1) { const long long now_ = clock64(); dbg_ph[i] += now_ - dbg_t1; dbg_t1 = now_; }
    for (int tick = 0; tick < nticks; ++tick) {
        const int slot = tick % R, nslot = (tick + 1) % R;
        if (p.debug) dbg_t0 = dbg_t1 = clock64();
        if ((tick % (R / 2)) == 0 && tick > 0) {
            // nobody may be more than half a ring behind before the slots written below are reused
            if (warp == 0) {
                const unsigned need = (unsigned)(tick - R / 2 + 1);
                unsigned n = 0;
                for (;;) {
                    unsigned mn = 0xFFFFFFFFu;
                    for (unsigned c = lane; c < gridDim.x; c += 32) mn = min(mn, *(volatile unsigned *)(p.progress + c));
                    for (int o = 16; o > 0; o >>= 1) mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o));
                    if (mn >= need) break;
                    if (++n > (1u << 20)) { if (lane == 0) atomicCAS(p.err, 0, 21); break; }
                    if (*(volatile int *)p.err != 0) break;
                }
            }
            __syncthreads();
        }
        if (is_lstm || is_proj) {
            if (is_lstm && l == 0 && lane < B && tick < len) {
                const int x = min(max(x_next, 0), p.V - 1);
                if (tick + 1 < len) x_next = __ldg(p.chars + o0 + tick + 1);
#pragma unroll
                for (int j = 0; j < KLM_MAX_UNITS_K; ++j) {
                    const int ui = warp + j * KLM_WARPS;
                    if (ui < role.count) {
                        const int u = role.first + ui;
                        pre[j] = __ldg(p.P + (size_t)x * W + u);      // (consumed only in the cell: no stall here)
                    }
                }
            }
            LL_DBG(0)                                          // progress check + gather issue
            // ---- stage the input vector(s): stage[b][K] <- words of this tick ----
            const int nsrc = (is_lstm && l > 0) ? 2 : 1;
            for (int src = 0; src < nsrc; ++src) {
                int hl;
                if (is_proj) hl = D - 1; else if (l == 0) hl = 0; else hl = (src == 0) ? l - 1 : l;
                const uint2 *hsrc = hx + (size_t)(slot * D + hl) * W * BS;
                for (int i = threadIdx.x; i < W * B; i += KLM_THREADS) {
                    const int k = i / B, b = i - k * B;
                    stage[b * K + src * W + k] = ll_wait(hsrc + (size_t)k * BS + b, (unsigned)tick, p.err);
                }
            }
            LL_DBG(1)                                          // own polls
            __syncthreads();
            LL_DBG(2)                                          // everybody's polls
            if (threadIdx.x == 0) __stcg(p.progress + blockIdx.x, (unsigned)(tick + 1));
            if (p.debug) { const long long now = clock64(); dbg_wait += now - dbg_t0; dbg_t0 = now; }
#pragma unroll
            for (int j = 0; j < KLM_MAX_UNITS_K; ++j) {
                const int ui = warp + j * KLM_WARPS;
                if (ui < role.count) {
                    float4 acc[BT];
#pragma unroll
                    for (int bb = 0; bb < BT; ++bb) acc[bb] = make_float4(0.f, 0.f, 0.f, 0.f);
                    const float4 *wrow = wptr + (size_t)ui * K;
#pragma unroll 4
                    for (int k = lane; k < K; k += 32) {
                        const float4 w = wrow[k];
#pragma unroll
                        for (int bb = 0; bb < BT; ++bb)
                            if (bb < B) fma4(acc[bb], w, stage[bb * K + k]);
                    }
                    float4 mine = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                    for (int bb = 0; bb < BT; ++bb) {
                        // 4 sums over 32 lanes in 6 shuffles: halve the set of values a lane carries while folding
                        const bool up16 = lane & 16, up8 = lane & 8;
                        float s0 = up16 ? acc[bb].x : acc[bb].z, s1 = up16 ? acc[bb].y : acc[bb].w;   // what the partner keeps
                        float k0 = up16 ? acc[bb].z : acc[bb].x, k1 = up16 ? acc[bb].w : acc[bb].y;   // what this lane keeps
                        k0 += __shfl_xor_sync(0xffffffffu, s0, 16);
                        k1 += __shfl_xor_sync(0xffffffffu, s1, 16);
                        float v = (up8 ? k1 : k0) + __shfl_xor_sync(0xffffffffu, up8 ? k0 : k1, 8);
                        v += __shfl_xor_sync(0xffffffffu, v, 4);
                        v += __shfl_xor_sync(0xffffffffu, v, 2);
                        v += __shfl_xor_sync(0xffffffffu, v, 1);
                        // lanes 0-7 hold x, 8-15 y, 16-23 z, 24-31 w: collect them in lane bb
                        const float gx = __shfl_sync(0xffffffffu, v, 0), gy = __shfl_sync(0xffffffffu, v, 8),
                                    gz = __shfl_sync(0xffffffffu, v, 16), gw = __shfl_sync(0xffffffffu, v, 24);
                        if (lane == bb) mine = make_float4(gx, gy, gz, gw);
                    }
                    LL_DBG(3)                                  // contraction + reduction
                    if (lane < B) {
                        const int u = role.first + ui, b = lane;
                        if (is_lstm) {
                            const int t = tick - l;
                            if (t >= 0 && t < len) {
                                add4(mine, pre[j]);
                                if (l == 0) {
                                    if (t < nreal) add4(mine, qsum[j]);
                                    else for (int k = 0; k < p.Nc; ++k) add4(mine, __ldg(p.Q[k] + u));   // padding: context 0
                                }
                                lstm_cell_fast(mine, c_reg[j], h_reg[j]);
                            }
                            outbuf[ui * B + b] = h_reg[j];
                        } else {
                            outbuf[(4 * ui) * B + b] = mine.x; outbuf[(4 * ui + 1) * B + b] = mine.y;
                            outbuf[(4 * ui + 2) * B + b] = mine.z; outbuf[(4 * ui + 3) * B + b] = mine.w;
                        }
                    }
                }
            }
            LL_DBG(4)                                          // cell
            __syncthreads();                                 // stage[] is rewritten by the next tick
            {
                // the CTA's words leave as ONE contiguous, coalesced run (whole sectors): single-lane 8-byte stores
                // scattered over the warps were measured to reach the pollers thousands of clocks later
                const int nw = (is_lstm ? role.count : 4 * role.count) * B;
                uint2 *dst = is_lstm ? hx + ((size_t)(nslot * D + l) * W + role.first) * B
                                     : lgx + ((size_t)nslot * 4 * p.V4 + 4 * role.first) * B;
                for (int i = threadIdx.x; i < nw; i += KLM_THREADS)
                    st_volatile_u2(dst + i, __float_as_uint(outbuf[i]), (unsigned)(tick + 1));
            }
            LL_DBG(5)                                          // barrier + publication
            if (p.debug) dbg_work += clock64() - dbg_t0;
        } else if (role.kind == KLM_ROLE_LSE) {
            // ---- log-sum-exp + target pick: one warp per stream, lanes over the vocabulary; ONE pass over the
            // words (running maximum and rescaled sum, the lane that meets the target keeps its logit), the target
            // index was fetched a tick ahead: the stage costs one L2 round trip per tick ----
            const int t = tick - D - 1;
            if (t >= 0) {
                const uint2 *lg = lgx + (size_t)slot * 4 * p.V4 * BS;
                for (int b = warp; b < B; b += KLM_WARPS) {
                    const long long ob = p.offsets[b];
                    const int lb = (int)(p.offsets[b + 1] - ob);
                    if (t >= lb) continue;
                    int tg = 0;
                    if (!p.full_rows) {
                        tg = (b < KLM_WARPS && tg_next_t == t) ? tg_next : __ldg(p.targets + ob + t);
                        tg = min(max(tg, 0), p.V - 1);
                        if (b < KLM_WARPS && t + 1 < lb) { tg_next = __ldg(p.targets + ob + t + 1); tg_next_t = t + 1; }
                    }
                    float m = -INFINITY, sum = 0.f, tl = 0.f;
                    for (int v0 = 0; v0 < p.V; v0 += 256) {
                        float x[8];
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const int v = v0 + 32 * j + lane;
                            x[j] = v < p.V ? ll_wait(lg + (size_t)v * BS + b, (unsigned)tick, p.err) : -INFINITY;
                            if (v == tg) tl = x[j];
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
                    if (p.debug && b == 0) { const long long now = clock64(); dbg_wait += now - dbg_t0; dbg_t0 = now; }
                    float mw = m;
                    for (int s = 16; s > 0; s >>= 1) mw = fmaxf(mw, __shfl_xor_sync(0xffffffffu, mw, s));
                    sum = (m > -INFINITY) ? sum * __expf(m - mw) : 0.f;
                    for (int s = 16; s > 0; s >>= 1) {
                        sum += __shfl_xor_sync(0xffffffffu, sum, s);
                        tl += __shfl_xor_sync(0xffffffffu, tl, s);       // exactly one lane holds the target's logit
                    }
                    if (p.full_rows) {
                        float *row = p.out + (size_t)(ob + t) * p.V;
                        const float inv = 1.0f / sum;
                        for (int v = lane; v < p.V; v += 32)
                            row[v] = expf(__uint_as_float(ld_volatile_u2(lg + (size_t)v * BS + b).x) - mw) * inv;
                    } else if (lane == 0) {
                        p.out[ob + t] = (tl - mw) - logf(sum);
                    }
                }
            }
            __syncthreads();
            if (threadIdx.x == 0) __stcg(p.progress + blockIdx.x, (unsigned)(tick + 1));
            if (p.debug) dbg_work += clock64() - dbg_t0;
        } else {
            if (threadIdx.x == 0) __stcg(p.progress + blockIdx.x, (unsigned)(tick + 1));
        }
    }

    if (p.debug > 1 && (threadIdx.x == 0 || threadIdx.x == 256) && (is_lstm || is_proj))
        printf("klm_seq_fma_ll cta %d kind %d layer %d thread %d: check+gather %lld, own polls %lld, all polls %lld, contraction %lld, cell %lld, "
               "sync+publish %lld clk per tick\n", blockIdx.x, role.kind, role.layer, threadIdx.x, dbg_ph[0] / nticks, dbg_ph[1] / nticks,
               dbg_ph[2] / nticks, dbg_ph[3] / nticks, dbg_ph[4] / nticks, dbg_ph[5] / nticks);
    if (p.debug && threadIdx.x == 0)
        printf("klm_seq_fma_ll cta %d kind %d layer %d first %d count %d: wait %lld work %lld clk per tick\n", blockIdx.x, role.kind,
               role.layer, role.first, role.count, dbg_wait / nticks, dbg_work / nticks);
    if (is_lstm) {
#pragma unroll
        for (int j = 0; j < KLM_MAX_UNITS_K; ++j) {
            const int ui = warp + j * KLM_WARPS;
            if (ui < role.count && lane < B) {
                const int u = role.first + ui;
                float *st = p.state + ((size_t)p.slots_out[lane] * D + l) * 2 * W;
                st[u] = h_reg[j]; st[W + u] = c_reg[j];
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Launch helper used by klm_api.cu
cudaError_t klm_launch_seq_fma(const KlmSeqParams &p, int grid, size_t smem, cudaStream_t stream) {
    void *args[] = { const_cast<KlmSeqParams *>(&p) };
    const void *fn;
    if (p.B > KLM_MODEK_MAXB) fn = (const void *)klm_seq_fma_streams;
    else if (p.hbuf_ll && p.B > 4) fn = (const void *)klm_seq_fma_ll<8>;
    else if (p.hbuf_ll && p.B > 2) fn = (const void *)klm_seq_fma_ll<4>;
    else if (p.hbuf_ll && p.B > 1) fn = (const void *)klm_seq_fma_ll<2>;
    else if (p.hbuf_ll) fn = (const void *)klm_seq_fma_ll<1>;
    else if (p.B > 4) fn = (const void *)klm_seq_fma_kpar<8>;
    else if (p.B > 2) fn = (const void *)klm_seq_fma_kpar<4>;
    else if (p.B > 1) fn = (const void *)klm_seq_fma_kpar<2>;
    else fn = (const void *)klm_seq_fma_kpar<1>;
    cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    return cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(KLM_THREADS), args, smem, stream);
}
