// "Wide" streamed-weights form of the tensor-core sequence kernel: up to 128 concurrent streams per launch for models
// whose fp16 hi/lo weight images do not fit shared memory (width 1024: cfg5 of BASELINE.json, 10 000 lines).
// Replaces the per-window Session.run of Rater.rate (rating.py:516) for 65..128 streams per launch.
//
// klm_seq_tcs (klm_seq_tc.cu) multiplies A = [64 streams hi ; 64 streams lo] (M = 128) with B = [N columns hi ; N lo]:
// four products per stream of which three are wanted, 2N accumulator columns, and every streamed B chunk serves 64
// streams.  Here a row of the accumulator is a STREAM and the three wanted products are accumulated into the SAME N
// columns by three MMAs per k-step:
//     D[128 x N] += h_hi . W_hi^T  +  h_hi . W_lo^T  +  h_lo . W_hi^T        (A tiles h_hi, h_lo: [W/8][128 rows][8] halves)
// -> 3 units of tensor work per 128 streams instead of 4 per 64, half the weight bytes per stream through the SM's
//    ingest port (the bound of the streamed form), N <= 192 TMEM columns, and NO hi/lo hand-over in the epilogue: a
//    thread owns a stream and half of the CTA's unit pairs, the pre-activation arrives complete.
//
// Same plan (roles, B images, scales) as klm_seq_tcs; same wavefront with skew 2 (REC l at 2l, IN l at 2l-1, PROJ at
// 2D-1, LSE at 2D), one grid barrier per tick.  Exchange buffers of its own (KlmTcParams::hx / Z / logits point at them):
//     hx     [D][2 parities][2 tiles: hi, lo][W/8][128][8] halves
//     Z      [D][2 slots][W units][128 streams] float4          logits [2 slots][Vp/4][128 streams] float4
// Shared memory: [1 KB barriers + bias of the slice][publication staging (REC): (pairs/4) x (2 KB hi + 2 KB lo)]
//                [ring of S <= 3 slots: A_hi chunk 16 KB | A_lo chunk 16 KB | B chunk 2 KB x pairs]
// Warp roles as in klm_seq_tc: warp 0 producer, warp 1 MMA issuer, warps 2..9 epilogue (row = 32 (warp & 3) + lane,
// the two warps of a TMEM lane quarter split the CTA's pairs).
#include <cstdio>
#include "klm_tc_dev.cuh"

// k per ring chunk: 64 (A tile chunk 16 KB) when three slots fit, else 32 with up to six slots (slices of 22 pairs)
#define TW_MAXR      6                      // epilogue rounds of 2 pairs per thread (<= 12 pairs per thread, 24 per CTA)
#define TW_HEAD      1024                   // barriers (first 128 bytes) + bias of the slice (from byte 256)

namespace {

struct TwCtx {
    uint32_t ring_a, bar_full, bar_empty, bar_tfull, bar_tempty, bar_go, tmem_base;
    uint32_t slot_bytes, cbb, ach;         // bytes of a ring slot, of its B part, of one A tile chunk
    int S, NCH, N, nticks, crot, src_layer, W, kc;
    unsigned G;
    const unsigned char *bsrc;
};

// Producer warp (converged; the elected lane issues): per tick the B parts of the first ring round go out before the
// grid barrier (they do not depend on the tick), the A tiles follow behind it.
template <bool TRACE>
__device__ __noinline__ void tw_produce(const KlmTcParams &p, const TwCtx &x) {
    const uint32_t ring_a = x.ring_a, bar_full = x.bar_full, bar_empty = x.bar_empty, bar_go = x.bar_go;
    const uint32_t slot_bytes = x.slot_bytes, cbb = x.cbb, ach = x.ach;
    const int S = x.S, NCH = x.NCH, nticks = x.nticks, crot = x.crot;
    const unsigned G = x.G;
    const unsigned char *bsrc = x.bsrc;
    const size_t tile_bytes = (size_t)x.W * 128 * sizeof(__half);                 // one tile of one parity of one layer
    const unsigned char *hx0 = reinterpret_cast<const unsigned char *>(p.hx) + (size_t)x.src_layer * 4 * tile_bytes;
    const bool lane0 = (threadIdx.x & 31) == 0;
    const int npre = S < NCH ? S : NCH;
    uint32_t phbits = 0;                                   // bit st: uses of slot st so far, mod 2
    for (int tick = 0; tick < nticks; ++tick) {
        for (int c0 = 0; c0 < npre; ++c0) {
            const int okw = tc_wait_mbar(bar_empty + 8 * c0, ((phbits >> c0) & 1u) ^ 1u, p, 2) ? 1 : 0;
            if (!__all_sync(0xffffffffu, okw)) return;
            if (tc::elect_one()) {
                const int cb = crot + c0 >= NCH ? crot + c0 - NCH : crot + c0;
                tc::mbar_arrive_expect_tx(bar_full + 8 * c0, 2 * ach + cbb);
                tc::bulk_g2s(ring_a + c0 * slot_bytes + 2 * ach, bsrc + (size_t)cb * cbb, cbb, bar_full + 8 * c0);
            }
            __syncwarp();
        }
        int okg = 1;
        if (lane0) {
            okg = tc_wait_grid(p, (unsigned)(tick + 1) * G, 1) ? 1 : 0;
            if (okg) tc::mbar_arrive(bar_go);
            if (TRACE && okg) tc_trace(p, tick, 0);
        }
        if (!__shfl_sync(0xffffffffu, okg, 0)) return;
        const unsigned char *src_hi = hx0 + (size_t)((tick & 1) * 2) * tile_bytes, *src_lo = src_hi + tile_bytes;
        int c = crot, st = 0;
        for (int c0 = 0; c0 < NCH; ++c0) {
            if (c0 >= npre) {
                const int okw = tc_wait_mbar(bar_empty + 8 * st, ((phbits >> st) & 1u) ^ 1u, p, 2) ? 1 : 0;
                if (!__all_sync(0xffffffffu, okw)) return;
            }
            if (tc::elect_one()) {
                const uint32_t dst = ring_a + st * slot_bytes, bar = bar_full + 8 * st;
                if (c0 >= npre) {
                    tc::mbar_arrive_expect_tx(bar, 2 * ach + cbb);
                    tc::bulk_g2s(dst + 2 * ach, bsrc + (size_t)c * cbb, cbb, bar);
                }
                tc::bulk_g2s(dst, src_hi + (size_t)c * ach, ach, bar);
                tc::bulk_g2s(dst + ach, src_lo + (size_t)c * ach, ach, bar);
                if (TRACE && c0 == 0) tc_trace(p, tick, 1);
            }
            __syncwarp();
            phbits ^= 1u << st;
            st = st + 1 == S ? 0 : st + 1;
            c = c + 1 == NCH ? 0 : c + 1;
        }
    }
}

// MMA warp: per k-step of 16 three MMAs into the same N accumulator columns (hi.hi, hi.lo, lo.hi).
template <bool TRACE>
__device__ __noinline__ void tw_mma(const KlmTcParams &p, const TwCtx &x) {
    const uint32_t idesc = tc::idesc_f16(128, x.N);
    const uint32_t lbo_a = 128 * 16, lbo_b = 2 * x.N * 16;
    const uint64_t a_base = tc::smem_desc(x.ring_a, lbo_a, 128);
    const uint64_t b_base = tc::smem_desc(x.ring_a + 2 * x.ach, lbo_b, 128);
    constexpr uint32_t a_k16 = (2 * 128 * 16) >> 4;
    const uint32_t a_lo = x.ach >> 4;
    const int nk16 = x.kc / 16;
    const uint32_t b_k16 = (2 * lbo_b) >> 4, b_lo = (uint32_t)x.N;           // lo rows start N * 16 bytes into a k group
    const uint32_t slot16 = x.slot_bytes >> 4;
    const uint32_t bar_full = x.bar_full, bar_empty = x.bar_empty, bar_tfull = x.bar_tfull, bar_tempty = x.bar_tempty;
    const uint32_t tmem_base = x.tmem_base;
    const int S = x.S, NCH = x.NCH, nticks = x.nticks;
    uint32_t phbits = 0;
    for (int tick = 0; tick < nticks; ++tick) {
        int okw = tc_wait_mbar(bar_tempty, (tick & 1) ^ 1, p, 3) ? 1 : 0;
        if (!__all_sync(0xffffffffu, okw)) return;
        tc::tc_fence_after();
        int st = 0;
        for (int c0 = 0; c0 < NCH; ++c0) {
            okw = tc_wait_mbar(bar_full + 8 * st, (phbits >> st) & 1u, p, 4) ? 1 : 0;
            if (!__all_sync(0xffffffffu, okw)) return;
            const uint64_t ad = a_base + (uint32_t)st * slot16, bd = b_base + (uint32_t)st * slot16;
            if (tc::elect_one()) {
                if (TRACE && (c0 == 0 || c0 == NCH - 1)) tc_trace(p, tick, c0 == 0 ? 2 : 3);
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) {
                    if (kk >= nk16) break;
                    const uint64_t a = ad + kk * a_k16, b = bd + kk * b_k16;
                    tc::mma_f16(tmem_base, a, b, idesc, (c0 | kk) != 0 ? 1u : 0u);        // h_hi . W_hi
                    tc::mma_f16(tmem_base, a, b + b_lo, idesc, 1u);                        // h_hi . W_lo
                    tc::mma_f16(tmem_base, a + a_lo, b, idesc, 1u);                        // h_lo . W_hi
                }
                tc::mma_commit(bar_empty + 8 * st);          // ring slot free once these MMAs have read it
                if (c0 == NCH - 1) tc::mma_commit(bar_tfull);   // accumulators of this tick complete
            }
            __syncwarp();
            phbits ^= 1u << st;
            st = st + 1 == S ? 0 : st + 1;
        }
    }
}

}  // namespace

namespace {

// Epilogue of one GEMM CTA.  KIND = KLM_TC_REC / KLM_TC_IN / KLM_TC_PROJ (compile-time: a REC CTA carries no IN state
// and vice versa -- the kernel runs at 168 registers per thread).  A thread owns stream s = 32 (warp & 3) + lane and NP / 2
// of the CTA's unit pairs, finished in rounds of two pairs (16 accumulator columns).
template <int KIND, bool TRACE>
__device__ __forceinline__ void tw_epilogue(const KlmTcParams &p, const KlmTcRole &role, const TwCtx &x, unsigned char *smem) {
    constexpr bool REC = KIND == KLM_TC_REC;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int W = p.W, l = role.layer, NP = role.npairs;
    const int q = warp & 3, s = 32 * q + lane, ch = (warp - 2) >> 2;
    const int NPT = NP / 2, pfirst = ch * NPT, nrounds = (NPT + 1) / 2;
    const uint32_t trow = x.tmem_base + ((uint32_t)(32 * q) << 16);
    const int sigma = REC ? 2 * l : KIND == KLM_TC_IN ? 2 * l - 1 : 2 * p.D - 1;
    const int nticks = x.nticks;
    const float scale = role.scale;
    const int ngrp = NP / 4;                                     // REC: groups of 8 units in the slice
    __half *pub = reinterpret_cast<__half *>(smem + TW_HEAD);    // [tile hi, lo][group][128 rows][8]
    float4 *bias_sm = reinterpret_cast<float4 *>(smem + 256);
    const size_t tile_halves = (size_t)W * 128;
    const bool pre0 = !(p.exp & 32);                             // (experiment switch: layer 1's table rows gathered per round)

    float c_reg[REC ? TW_MAXR : 1][2][2], h_reg[REC ? TW_MAXR : 1][2][2];
    // The row's queue of texts (continuous batching: KlmTcParams::lane_ptr): text `cur` occupies positions g0 .. g0 + len - 1
    // of the row's time line, position g0 + len is a gap tick in which the recurrence stages swap the state (final state of
    // the text to its slot, initial state of the next text in, and THAT is what they publish), then the next text starts.
    int cur = 0, qend = 0, sid = -1, len = 0, g0 = 0;
    long long o0 = 0;
    int x_next = 0, nreal_s = 0x7fffffff;
    int zk[KLM_MAX_CTX];
#pragma unroll
    for (int k = 0; k < KLM_MAX_CTX; ++k) zk[k] = 0;
    auto open_text = [&]() {
        sid = -1; len = 0; o0 = 0;
        if (cur < qend) { sid = __ldg(p.lane_streams + cur); o0 = p.offsets[sid]; len = (int)(p.offsets[sid + 1] - o0); }
    };
    // state of the open text in (zeros if the row has run out of texts); layer 1 also fetches what its input needs
    auto load_text_state = [&]() {
        const int slot = sid >= 0 ? p.slots_in[sid] : -1;
#pragma unroll
        for (int r = 0; r < (REC ? TW_MAXR : 0); ++r)
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                c_reg[r][i][0] = c_reg[r][i][1] = h_reg[r][i][0] = h_reg[r][i][1] = 0.f;
                if (2 * r + i < NPT && slot >= 0) {
                    const int u = 2 * (role.first_pair + pfirst + 2 * r + i);
                    const float *st = p.state + ((size_t)slot * p.D + l) * 2 * W;
                    const float2 h2 = *reinterpret_cast<const float2 *>(st + u), c2 = *reinterpret_cast<const float2 *>(st + W + u);
                    h_reg[r][i][0] = h2.x; h_reg[r][i][1] = h2.y; c_reg[r][i][0] = c2.x; c_reg[r][i][1] = c2.y;
                }
            }
        if (l == 0) {
            x_next = 0; nreal_s = 0x7fffffff;
            if (sid >= 0) {
                if (len > 0) x_next = __ldg(p.chars + o0);
                if (p.n_real) nreal_s = __ldg(p.n_real + sid);
#pragma unroll
                for (int k = 0; k < KLM_MAX_CTX; ++k)
                    if (k < p.Nc) zk[k] = min(max(__ldg(p.ctx + sid * p.Nc + k), 0), p.ctx_rows[k] - 1);
            }
        }
    };
    auto store_text_state = [&]() {
        if (sid < 0) return;
        float *st = p.state + ((size_t)p.slots_out[sid] * p.D + l) * 2 * W;
#pragma unroll
        for (int r = 0; r < (REC ? TW_MAXR : 0); ++r)
#pragma unroll
            for (int i = 0; i < 2; ++i)
                if (2 * r + i < NPT) {
                    const int u = 2 * (role.first_pair + pfirst + 2 * r + i);
                    *reinterpret_cast<float2 *>(st + u) = make_float2(h_reg[r][i][0], h_reg[r][i][1]);
                    *reinterpret_cast<float2 *>(st + W + u) = make_float2(c_reg[r][i][0], c_reg[r][i][1]);
                }
    };
    if constexpr (REC) {
        // ---- first text of the row: carried state in, initial hidden vector published for tick 0 (parity 0) ----
        cur = __ldg(p.lane_ptr + s); qend = __ldg(p.lane_ptr + s + 1);
        open_text();
        load_text_state();
#pragma unroll
        for (int r = 0; r < TW_MAXR; ++r)
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                if (2 * r + i < NPT) {
                    const int u = 2 * (role.first_pair + pfirst + 2 * r + i);
                    __half hi0, lo0, hi1, lo1;
                    tc_split(h_reg[r][i][0], hi0, lo0); tc_split(h_reg[r][i][1], hi1, lo1);
                    __half *dst = p.hx + (size_t)(l * 4) * tile_halves + (size_t)(u / 8) * (128 * 8) + (size_t)s * 8 + (u & 7);
                    *reinterpret_cast<__half2 *>(dst) = __halves2half2(hi0, hi1);
                    *reinterpret_cast<__half2 *>(dst + tile_halves) = __halves2half2(lo0, lo1);
                }
            }
        tc::fence_proxy_async();                    // generic global writes -> later bulk (async proxy) reads
    }
    if (KIND == KLM_TC_IN)
        for (int j = tid - 64; j < 2 * NP; j += 256) bias_sm[j] = __ldg(p.bias[l] + 2 * role.first_pair + j);
    // everything this CTA must have written before anybody's first fetch is out: first arrival at the grid barrier
    asm volatile("bar.sync 2, 320;" ::: "memory");
    if (tid == 64) tc::red_release_add(p.barrier, 1u);

    for (int tick = 0; tick < nticks; ++tick) {
        int ok = tc_wait_mbar(x.bar_go, tick & 1, p, 5) ? 1 : 0;
        if (!__all_sync(0xffffffffu, ok)) break;
        int t = tick - sigma - g0;
        if (REC && sid >= 0 && t == len) {
            // gap tick behind text `cur`: its final state out, the next text's initial state in (its loads are in flight while
            // the operand streams in; the rounds below stage and publish the new hidden values unchanged: nothing is active)
            store_text_state();
            ++cur; g0 += len + 1;
            open_text();
            load_text_state();
            t = -1;
        }
        const bool active = REC && sid >= 0 && t >= 0 && t < len;
        const int par_out = (tick + 1) & 1, zs_in = tick & 1, zs_out = (tick + 1) & 1;
        int xc = 0;
        bool real = true;
        if (REC && l == 0 && active) {
            xc = min(max(x_next, 0), p.V - 1);
            if (t + 1 < len) x_next = __ldg(p.chars + o0 + t + 1);
            real = t < nreal_s;
        }
        // addends of round r (REC): layer 1 = fused char + context embedding gather (tables P, Q: rating.py:103-115,125;
        // padding positions take context 0, :1141-1143); layers above = Z of the IN stage
        auto load_add = [&](const int r, float4 (&a4)[2][2]) {
#pragma unroll
            for (int i = 0; i < 2; ++i) a4[i][0] = a4[i][1] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (!REC || !active) return;
#pragma unroll
            for (int i = 0; i < 2; ++i)
                if (2 * r + i < NPT) {
                    const int u = 2 * (role.first_pair + pfirst + 2 * r + i);
                    if (l == 0 && (r == 0 || !pre0)) {
                        float4 a0, a1;
                        tc_ldg_pair(p.P + (size_t)xc * W + u, a0, a1);
#pragma unroll
                        for (int k = 0; k < KLM_MAX_CTX; ++k)
                            if (k < p.Nc) {
                                float4 b0, b1;
                                tc_ldg_pair(p.Q[k] + (real ? (size_t)zk[k] * W : (size_t)0) + u, b0, b1);
                                a0.x += b0.x; a0.y += b0.y; a0.z += b0.z; a0.w += b0.w;
                                a1.x += b1.x; a1.y += b1.y; a1.z += b1.z; a1.w += b1.w;
                            }
                        a4[i][0] = a0; a4[i][1] = a1;
                    } else {
                        // layers above: Z of the IN stage; layer 1, rounds >= 1: what this thread gathered at the top of the tick
                        const float4 *zrow = reinterpret_cast<const float4 *>(p.Z) + ((size_t)(l * 2 + zs_in) * W) * 128 + s;
                        a4[i][0] = __ldcg(zrow + (size_t)u * 128); a4[i][1] = __ldcg(zrow + (size_t)(u + 1) * 128);
                    }
                }
        };
        float4 add4[2][2];
        load_add(0, add4);
        if (REC && pre0 && l == 0 && active) {
            // Layer 1, rounds >= 1: the table rows are gathered NOW, while the operand streams in and this warp has nothing
            // to do, and parked in the layer's own (otherwise unused) Z slot in the coalesced [unit][128 streams] layout; the
            // rounds then fetch them a round ahead like the layers above fetch Z.  Gathered just in time they cost an exposed
            // L2 round trip per round (~3 000 clk of the tick: REC l0 was the last stage to reach the grid barrier).
            float4 *zrow = reinterpret_cast<float4 *>(p.Z) + ((size_t)zs_in * W) * 128 + s;
            for (int r = 1; r < nrounds; ++r) {
#pragma unroll
                for (int i = 0; i < 2; ++i)
                    if (2 * r + i < NPT) {
                        const int u = 2 * (role.first_pair + pfirst + 2 * r + i);
                        float4 a0, a1;
                        tc_ldg_pair(p.P + (size_t)xc * W + u, a0, a1);
#pragma unroll
                        for (int k = 0; k < KLM_MAX_CTX; ++k)
                            if (k < p.Nc) {
                                float4 b0, b1;
                                tc_ldg_pair(p.Q[k] + (real ? (size_t)zk[k] * W : (size_t)0) + u, b0, b1);
                                a0.x += b0.x; a0.y += b0.y; a0.z += b0.z; a0.w += b0.w;
                                a1.x += b1.x; a1.y += b1.y; a1.z += b1.z; a1.w += b1.w;
                            }
                        __stcg(zrow + (size_t)u * 128, a0); __stcg(zrow + (size_t)(u + 1) * 128, a1);
                    }
            }
        }

        ok = tc_wait_mbar(x.bar_tfull, tick & 1, p, 6) ? 1 : 0;
        if (!__all_sync(0xffffffffu, ok)) break;
        if (TRACE && tid == 64) tc_trace(p, tick, 4);
        tc::tc_fence_after();
#pragma unroll
        for (int r = 0; r < TW_MAXR; ++r) {
            if (r >= nrounds) break;                            // (CTA-uniform)
            float acc[2][8];
#pragma unroll
            for (int i = 0; i < 2; ++i) {
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
                if (2 * r + i < NPT) tc::tmem_ld8(trow + 8 * (pfirst + 2 * r + i), acc[i]);
            }
            tc::tmem_ld_wait();
            if (r == nrounds - 1) {
                tc::tc_fence_before();
                tc::mbar_arrive(x.bar_tempty);                  // this thread is done reading TMEM
                if (TRACE && tid == 64) tc_trace(p, tick, 5);
            }
            float4 add_nxt[2][2];
            if (REC && (l > 0 || pre0) && r + 1 < nrounds) load_add(r + 1, add_nxt);      // (coalesced: cheap to issue, in flight under the cell)
            if constexpr (REC) {
                // LSTM cell (Keras LSTMCell, gate order i|f|c|o, exact logistic: rating.py:133-134); 8 MUFU per unit, staged
                float e1[2][2], e2[2][2], e3[2][2], e4[2][2], ag[2][2];
#pragma unroll
                for (int i = 0; i < 2; ++i)
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const float4 a = add4[i][e];
                        ag[i][e] = fmaf(acc[i][4 * e + 2], scale, a.z);
                        e1[i][e] = tc_exp(-tc_clamp(fmaf(acc[i][4 * e], scale, a.x)));
                        e2[i][e] = tc_exp(-tc_clamp(fmaf(acc[i][4 * e + 1], scale, a.y)));
                        e3[i][e] = tc_exp(-2.f * tc_clamp(ag[i][e]));
                        e4[i][e] = tc_exp(-tc_clamp(fmaf(acc[i][4 * e + 3], scale, a.w)));
                    }
#pragma unroll
                for (int i = 0; i < 2; ++i)
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        float num, den;
                        klm_tanh_parts(ag[i][e], e3[i][e], num, den);
                        const float ig = __fdividef(num, (1.f + e1[i][e]) * den);
                        const float f = __fdividef(1.f, 1.f + e2[i][e]);
                        e1[i][e] = fmaf(f, c_reg[r][i][e], ig);                    // new cell state
                        e2[i][e] = tc_exp(-2.f * tc_clamp(e1[i][e]));
                    }
#pragma unroll
                for (int i = 0; i < 2; ++i) {
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        float num, den;
                        klm_tanh_parts(e1[i][e], e2[i][e], num, den);
                        const float hn = __fdividef(num, (1.f + e4[i][e]) * den);
                        if (active) { c_reg[r][i][e] = e1[i][e]; h_reg[r][i][e] = hn; }
                    }
                    if (2 * r + i < NPT) {
                        // the new hidden values as fp16 hi / lo in the readers' operand layout
                        const int ul = 2 * (pfirst + 2 * r + i);
                        __half hi0, lo0, hi1, lo1;
                        tc_split(h_reg[r][i][0], hi0, lo0); tc_split(h_reg[r][i][1], hi1, lo1);
                        __half *dst = pub + ((size_t)(ul / 8) * 128 + s) * 8 + (ul & 7);
                        *reinterpret_cast<__half2 *>(dst) = __halves2half2(hi0, hi1);
                        *reinterpret_cast<__half2 *>(dst + (size_t)ngrp * 128 * 8) = __halves2half2(lo0, lo1);
                    }
                }
                if (r + 1 < nrounds) {
                    if (l > 0 || pre0) {
#pragma unroll
                        for (int i = 0; i < 2; ++i) { add4[i][0] = add_nxt[i][0]; add4[i][1] = add_nxt[i][1]; }
                    } else {
                        load_add(r + 1, add4);      // (scattered table rows: issued behind the cell, see klm_seq_tc.cu)
                    }
                }
            } else if constexpr (KIND == KLM_TC_IN) {
                float4 *zrow = reinterpret_cast<float4 *>(p.Z) + ((size_t)(l * 2 + zs_out) * W) * 128 + s;
#pragma unroll
                for (int i = 0; i < 2; ++i)
                    if (2 * r + i < NPT) {
                        const int pl = pfirst + 2 * r + i, u = 2 * (role.first_pair + pl);
#pragma unroll
                        for (int e = 0; e < 2; ++e) {
                            const float4 b = bias_sm[2 * pl + e];
                            __stcg(zrow + (size_t)(u + e) * 128,
                                   make_float4(fmaf(acc[i][4 * e], scale, b.x), fmaf(acc[i][4 * e + 1], scale, b.y),
                                               fmaf(acc[i][4 * e + 2], scale, b.z), fmaf(acc[i][4 * e + 3], scale, b.w)));
                        }
                    }
            } else {   // PROJ: logits of vocabulary rows 8 * (first_pair + pair) ..
                float4 *lrow = reinterpret_cast<float4 *>(p.logits) + ((size_t)zs_out * (p.Vp / 4)) * 128 + s;
#pragma unroll
                for (int i = 0; i < 2; ++i)
                    if (2 * r + i < NPT) {
                        const int g8 = role.first_pair + pfirst + 2 * r + i;
                        __stcg(lrow + (size_t)(2 * g8) * 128, make_float4(acc[i][0] * scale, acc[i][1] * scale, acc[i][2] * scale, acc[i][3] * scale));
                        __stcg(lrow + (size_t)(2 * g8 + 1) * 128, make_float4(acc[i][4] * scale, acc[i][5] * scale, acc[i][6] * scale, acc[i][7] * scale));
                    }
            }
        }
        // ---- hand over ----
        if (REC) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");       // staging writes -> bulk store reads
        if (TRACE && tid == 64) tc_trace(p, tick, 6);
        tc_bar_epi();
        if (tid == 64) {
            if (REC) {
                __half *ghi = p.hx + (size_t)((l * 2 + par_out) * 2) * tile_halves + (size_t)(role.first_pair / 4) * (128 * 8);
                const uint32_t bytes = (uint32_t)ngrp * 2048u;
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                             ::"l"(ghi), "r"(tc::smem_u32(pub)), "r"(bytes) : "memory");
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                             ::"l"(ghi + tile_halves), "r"(tc::smem_u32(pub) + bytes), "r"(bytes) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");       // writes complete (not just read)
            }
            tc::red_release_add(p.barrier, 1u);
            if (TRACE) tc_trace(p, tick, 7);
        }
        // (the staging is rewritten only behind the next tick's grid barrier, which this CTA's arrival -- behind the completed
        //  bulk store -- is part of)
    }
    // (the final state of a row's last text went out in its gap tick: the launch covers it; a text still open here means the
    //  host sized the launch too short)
    if constexpr (REC) store_text_state();
}

}  // namespace


// LSE stage of the wide form: tc_lse_role (klm_tc_dev.cuh) with rows of 128 streams and the per-row text queues -- lane =
// row, a warp covers 32 rows and one slice of the vocabulary; the row's text and position follow from the tick as in the
// epilogue of the recurrence stages (a gap tick between two texts).
static __device__ void tw_lse_role(const KlmTcParams &p, const KlmTcRole &role, unsigned char *dyn_smem) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const unsigned G = gridDim.x;
    const int nticks = p.T_max + 2 * p.D;
    const int V = p.V, Vq = p.Vp / 4, V4 = (V + 3) / 4, sigma = 2 * p.D, full_rows = p.full_rows;
    const float4 *logits4 = reinterpret_cast<const float4 *>(p.logits);
    const int *targets = p.targets;
    float *out = p.out;
    __shared__ int ok_s;
    float (*part_m)[32] = reinterpret_cast<float (*)[32]>(dyn_smem);
    float (*part_s)[32] = reinterpret_cast<float (*)[32]>(dyn_smem + TC_LSE_WARPS * 32 * sizeof(float));
    const int ngroups = (p.B + 31) / 32;                                     // p.B = rows in use
    const int gmine = role.first_pair < ngroups ? (ngroups - role.first_pair + role.npairs - 1) / role.npairs : 0;
    const int nslices = gmine > 0 ? TC_LSE_WARPS / gmine : 1;
    const int gl = warp / nslices, sl = warp - gl * nslices;
    const bool warp_on = gmine > 0 && gl < gmine;
    const int s = warp_on ? 32 * (role.first_pair + gl * role.npairs) + lane : 0;
    int cur = 0, qend = 0, sid = -1, len = 0, g0 = 0;
    long long o0 = 0;
    auto open_text = [&]() {
        sid = -1; len = 0; o0 = 0;
        if (cur < qend) { sid = __ldg(p.lane_streams + cur); o0 = p.offsets[sid]; len = (int)(p.offsets[sid + 1] - o0); }
    };
    if (warp_on && s < 128) { cur = __ldg(p.lane_ptr + s); qend = __ldg(p.lane_ptr + s + 1); open_text(); }
    const int v4_lo = (int)((long long)V4 * sl / nslices), v4_hi = (int)((long long)V4 * (sl + 1) / nslices);
    if (tid == 0) tc::red_release_add(p.barrier, 1u);
    for (int tick = 0; tick < nticks; ++tick) {
        if (tid == 0) ok_s = tc_wait_grid(p, (unsigned)(tick + 1) * G, 10) ? 1 : 0;
        __syncthreads();
        if (!ok_s) return;
        int t = tick - sigma - g0;
        if (sid >= 0 && t == len) { ++cur; g0 += len + 1; open_text(); t = -1; }      // gap tick: next text of the row
        const bool act = warp_on && sid >= 0 && t >= 0 && t < len;
        const float4 *base = logits4 + (size_t)(tick & 1) * Vq * 128 + s;           // (slot written by PROJ one tick ago)
        float tl = 0.f;
        if (act && sl == 0 && !full_rows) {
            int tg = __ldg(targets + o0 + t);
            tg = min(max(tg, 0), V - 1);
            tl = __ldcg(reinterpret_cast<const float *>(base + (size_t)(tg >> 2) * 128) + (tg & 3));
        }
        float m = -INFINITY, sum = 0.f;
        if (warp_on && tick >= sigma) {
            for (int v0 = v4_lo; v0 < v4_hi; v0 += 8) {
                float4 x[8];
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    x[j] = (v0 + j < v4_hi) ? __ldcg(base + (size_t)(v0 + j) * 128) : make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
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
                    const float4 x = __ldcg(base + (size_t)v4 * 128);
                    const int v = 4 * v4;
                    dst[v] = expf(x.x - M) * inv;
                    if (v + 1 < V) dst[v + 1] = expf(x.y - M) * inv;
                    if (v + 2 < V) dst[v + 2] = expf(x.z - M) * inv;
                    if (v + 3 < V) dst[v + 3] = expf(x.w - M) * inv;
                }
            }
        }
        __syncthreads();
        if (tid == 0) tc::red_release_add(p.barrier, 1u);
    }
}

template <bool TRACE>
static __device__ __forceinline__ void tw_body(const KlmTcParams &p, unsigned char *smem) {
    const KlmTcRole role = p.roles[blockIdx.x];
    if (role.kind == KLM_TC_LSE) { tw_lse_role(p, role, smem); return; }

    const int tid = threadIdx.x, warp = tid >> 5;
    const int NP = role.npairs, N = 8 * NP, W = p.W;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem);
    const uint32_t bar_full = tc::smem_u32(bars), bar_empty = tc::smem_u32(bars + 6);            // up to 6 ring slots
    const uint32_t bar_tfull = tc::smem_u32(bars + 12), bar_tempty = tc::smem_u32(bars + 13), bar_go = tc::smem_u32(bars + 14);
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 15);
    // ring geometry: everything behind the head and (REC) the publication staging, at most 3 slots
    const uint32_t pub_bytes = role.kind == KLM_TC_REC ? (uint32_t)(NP / 4) * 4096u : 0u;
    const uint32_t room = (uint32_t)p.ring_bytes - TW_HEAD - pub_bytes;
    int kc = 64;
    if (3u * (2u * 64u * 256u + 64u * 32u * (uint32_t)NP) > room && !(p.exp & 64)) kc = 32;      // three slots of 64 k do not fit: finer chunks (+3 % on cfg5 over two slots of 64 k)
    const uint32_t ach = (uint32_t)kc * 256u;                             // one chunk of one A tile: kc k x 128 rows x 2 B
    const uint32_t cbb = (uint32_t)kc * 32u * (uint32_t)NP;               // kc / 8 k groups x 2N rows x 16 bytes
    const uint32_t slot_bytes = 2u * ach + cbb;
    const int S = min(kc == 64 ? 3 : 6, (int)(room / slot_bytes));
    if (S < 2 || NP < 2 || (NP & 1) || NP > 4 * TW_MAXR || (role.kind == KLM_TC_REC && (NP & 3))) {
        if (tid == 0) atomicCAS(p.err, 0, 91);               // (the host checks the plan before it launches this kernel)
        return;
    }
    if (warp == 0) {
        tc::tmem_alloc(tc::smem_u32(tmem_slot), 256);
        tc::tmem_relinquish();
    }
    if (tid == 32) {
        for (int i = 0; i < S; ++i) { tc::mbar_init(bar_full + 8 * i, 1); tc::mbar_init(bar_empty + 8 * i, 1); }
        tc::mbar_init(bar_tfull, 1);
        tc::mbar_init(bar_tempty, 256);
        tc::mbar_init(bar_go, 1);
        tc::fence_mbar_init();
    }
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();

    TwCtx x;
    x.ring_a = tc::smem_u32(smem + TW_HEAD + pub_bytes);
    x.bar_full = bar_full; x.bar_empty = bar_empty; x.bar_tfull = bar_tfull; x.bar_tempty = bar_tempty; x.bar_go = bar_go;
    x.tmem_base = *tmem_slot;
    x.slot_bytes = slot_bytes; x.cbb = cbb; x.ach = ach; x.kc = kc;
    x.S = S; x.NCH = W / kc; x.N = N; x.nticks = p.T_max + 2 * p.D;
    x.crot = role.rot % (W / kc); x.src_layer = role.src_layer; x.W = W;
    x.G = gridDim.x;
    x.bsrc = p.wimg + role.wofs;

    if (warp == 0) {
        asm volatile("bar.sync 2, 320;" ::: "memory");       // (the epilogue warps' prologue: carried state in, bias staged)
        tw_produce<TRACE>(p, x);
    } else if (warp == 1) {
        asm volatile("bar.sync 2, 320;" ::: "memory");
        tw_mma<TRACE>(p, x);
    } else if (role.kind == KLM_TC_REC) {
        tw_epilogue<KLM_TC_REC, TRACE>(p, role, x, smem);
    } else if (role.kind == KLM_TC_IN) {
        tw_epilogue<KLM_TC_IN, TRACE>(p, role, x, smem);
    } else {
        tw_epilogue<KLM_TC_PROJ, TRACE>(p, role, x, smem);
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        __syncwarp();
        tc::tmem_dealloc(x.tmem_base, 256);
    }
}

__global__ void __launch_bounds__(KLM_TC_THREADS, 1) klm_seq_tw(const __grid_constant__ KlmTcParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    tw_body<false>(p, smem);
}
// timeline instrumentation compiled in: launched only while klm_tc_trace is armed
__global__ void __launch_bounds__(KLM_TC_THREADS, 1) klm_seq_tw_traced(const __grid_constant__ KlmTcParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    tw_body<true>(p, smem);
}

// ring_bytes of the parameters = dynamic shared memory of the launch (the kernel derives the ring geometry from it)
cudaError_t klm_launch_seq_tw(const KlmTcParams &p, int grid, size_t smem, cudaStream_t stream) {
    void *args[] = { const_cast<KlmTcParams *>(&p) };
    static size_t smem_set[2][64] = {{0}};
    const int which = p.trace ? 1 : 0;
    const void *fn = which ? (const void *)klm_seq_tw_traced : (const void *)klm_seq_tw;
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64 || smem_set[which][dev] != smem) {
        cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        if (dev >= 0 && dev < 64) smem_set[which][dev] = smem;
    }
    return cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(KLM_TC_THREADS), args, smem, stream);
}
