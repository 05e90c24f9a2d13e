// Persistent stacked-LSTM sequence kernel on the 5th-generation tensor cores (tcgen05 + TMEM),
// for 9..64 concurrent streams / hypotheses.
//
// Replaces the per-window Session.run of Rater.rate (rating.py:516) and the per-step one of
// Rater.predict (:631) when many streams advance together (rate_many, beams).
//
// Every CTA owns, for the whole launch, N <= 80 columns of ONE weight matrix, resident in shared
// memory as an fp16 hi/lo split in the tcgen05 operand layout ("B image", [W/8][2N][8] halves,
// rows 0..N-1 = hi, N..2N-1 = lo):
//   REC l : recurrent kernel R_l of LSTM layer l          (gate columns unit-interleaved i,f,c,o)
//   IN  l : input kernel K_l of layer l >= 1               (layer 1's input is the table P, no GEMM)
//   PROJ  : rows of the char embedding E (tied output projection)
//   LSE   : no weights; log-sum-exp over the logits, target pick / full softmax rows
// Per tick a GEMM CTA computes D[128 x 2N] = A[128 x W] . B[2N x W]^T where A is one layer's
// hidden vector for all 64 streams, rows 0..63 the fp16 hi part and rows 64..127 the lo part,
// fetched from the exchange buffer `hx` with bulk async copies through an mbarrier ring.
// pre-activation = D[s][j] + D[s][N+j] + D[64+s][j]   (hi.hi + hi.lo + lo.hi; lo.lo is dropped)
//
// Wavefront over ticks, stage sigma works on position t = tick - sigma:
//   REC l: sigma = 2l, IN l: 2l-1, PROJ: 2D-1, LSE: 2D.   IN l writes Z_l[t] (+ bias), REC l adds it
// one tick later, so each CTA needs exactly one hidden vector per tick.  One grid-wide barrier per
// tick (release/acquire counter); everything handed over is double-buffered by tick parity.
//
// Warp roles in a GEMM CTA (192 threads): warp 0 lane 0 = producer (grid barrier wait + bulk
// copies), warp 1 lane 0 = MMA issuer, warps 2..5 = epilogue (TMEM -> registers, hi/lo halves
// exchanged through shared memory, LSTM cell in registers, hand-over stores).
#include "klm_tc_dev.cuh"

// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(KLM_TC_THREADS, 1) klm_seq_tc(const KlmTcParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    const KlmTcRole role = p.roles[blockIdx.x];
    if (role.kind == KLM_TC_LSE) { tc_lse_role(p, role); return; }

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int W = p.W, S = p.stages;
    const unsigned G = gridDim.x;
    const int nticks = p.T_max + 2 * p.D;
    const int NP = role.npairs, N = 8 * NP;
    const int nchunks = W / KLM_TC_KC;

    // ---- shared memory carve-up ----
    unsigned char *ring = smem;                                          // S x 16 KB: A operand chunks
    float *xchg = reinterpret_cast<float *>(smem + KLM_TC_XCHG_OFS);    // 20 KB hi<->lo hand-over; aliases the ring when
                                                                         // S >= 2 (the ring is idle while the epilogue runs)
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + KLM_TC_FIXED_BYTES(S) - 128);
    unsigned char *wsm = smem + KLM_TC_FIXED_BYTES(S);                   // B image
    const uint32_t bar_full = tc::smem_u32(bars), bar_empty = tc::smem_u32(bars + 4);
    const uint32_t bar_tfull = tc::smem_u32(bars + 8), bar_tempty = tc::smem_u32(bars + 9);
    const uint32_t bar_go = tc::smem_u32(bars + 10);       // producer -> epilogue: grid barrier of this tick passed
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 11);

    if (warp == 0) {
        tc::tmem_alloc(tc::smem_u32(tmem_slot), KLM_TC_TMEM_COLS);
        tc::tmem_relinquish();
    }
    if (tid == 32) {
        for (int i = 0; i < S; ++i) { tc::mbar_init(bar_full + 8 * i, p.ingest_bulk ? 1 : 256); tc::mbar_init(bar_empty + 8 * i, 1); }
        tc::mbar_init(bar_tfull, 1);
        tc::mbar_init(bar_tempty, 256);
        tc::mbar_init(bar_go, 1);
        tc::fence_mbar_init();
    }
    {   // B image: global -> shared, once per launch
        const uint4 *src = reinterpret_cast<const uint4 *>(p.wimg + role.wofs);
        uint4 *dst = reinterpret_cast<uint4 *>(wsm);
        const int n16 = (W / 8) * 2 * N;
        for (int i = tid; i < n16; i += KLM_TC_THREADS) dst[i] = __ldg(src + i);
        tc::fence_proxy_async();                     // generic writes -> visible to the MMA's operand reads
    }

    // ---- epilogue thread identity ----
    const int q = warp & 3;                          // TMEM lane quarter this warp may read
    const int row = 32 * q + lane;                   // accumulator row: 0..63 hi part, 64..127 lo part
    const int half = row >> 6, s = row & 63;
    const int ch = (warp - 2) >> 2;                  // two warps share a lane quarter and split its pairs
    const int Ph = (NP + 1) / 2;                     // pairs [0,Ph) are finished by hi-row threads, the rest by lo-row threads
    const int my_first = half ? Ph : 0, my_n = half ? NP - Ph : Ph;          // this half's pairs
    const int oth_first = half ? 0 : Ph, oth_n = half ? Ph : NP - Ph;        // the other half's pairs
    const int own_first = my_first + (ch ? (my_n + 1) / 2 : 0), nown = ch ? my_n - (my_n + 1) / 2 : (my_n + 1) / 2;
    const int send_first = oth_first + (ch ? (oth_n + 1) / 2 : 0), nsend = ch ? oth_n - (oth_n + 1) / 2 : (oth_n + 1) / 2;
    const bool is_epi = warp >= 2;
    const bool is_rec = role.kind == KLM_TC_REC;
    const int l = role.layer;

    float c_reg[KLM_TC_OWN][2], h_reg[KLM_TC_OWN][2];
    long long o0 = 0;
    int len = 0;
    if (is_epi) {
        if (s < p.B) { o0 = p.offsets[s]; len = (int)(p.offsets[s + 1] - o0); }
#pragma unroll
        for (int i = 0; i < KLM_TC_OWN; ++i) {
            c_reg[i][0] = c_reg[i][1] = h_reg[i][0] = h_reg[i][1] = 0.f;
            if (is_rec && i < nown) {
                const int u = 2 * (role.first_pair + own_first + i);
                if (s < p.B) {
                    const int slot = p.slots_in[s];
                    if (slot >= 0) {
                        const float *st = p.state + ((size_t)slot * p.D + l) * 2 * W;
                        h_reg[i][0] = st[u]; h_reg[i][1] = st[u + 1];
                        c_reg[i][0] = st[W + u]; c_reg[i][1] = st[W + u + 1];
                    }
                }
                // publish the carried hidden state for tick 0 (parity 0)
                __half hi0, lo0, hi1, lo1;
                tc_split(h_reg[i][0], hi0, lo0); tc_split(h_reg[i][1], hi1, lo1);
                __half *dst = p.hx + ((size_t)(l * 2 + 0) * (W / 8) + u / 8) * (128 * 8) + (u & 7);
                *reinterpret_cast<__half2 *>(dst + s * 8) = __halves2half2(hi0, hi1);
                *reinterpret_cast<__half2 *>(dst + (64 + s) * 8) = __halves2half2(lo0, lo1);
            }
        }
        if (p.ingest_bulk) tc::fence_proxy_async(); // generic global writes -> later bulk (async proxy) reads
    }
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (tid == 0) tc::red_release_add(p.barrier, 1u);

    if (warp == 0) {
        // =============================== producer ===============================
        if (lane == 0) {
            uint32_t it = 0;
            bool ok = true;
            for (int tick = 0; tick < nticks && ok; ++tick) {
                if (!tc_wait_grid(p, (unsigned)(tick + 1) * G, 1)) break;
                tc::mbar_arrive(bar_go);
                tc_trace(p, tick, 0);
                if (p.ingest_bulk) {
                    // operand fetch by the bulk-copy engine (measured: ~25 B/clk per SM + ~90 clk per copy,
                    // slower than the load/store path below; kept selectable for comparison)
                    tc::fence_proxy_async();             // generic global writes of other CTAs -> async-proxy reads
                    const unsigned char *src = reinterpret_cast<const unsigned char *>(
                        p.hx + (size_t)(role.src_layer * 2 + (tick & 1)) * W * 128);
                    for (int c0 = 0; c0 < nchunks; ++c0, ++it) {
                        const int c = (c0 + (int)blockIdx.x) % nchunks;    // spread the CTAs over the chunks
                        const uint32_t st = it % S, ph = (it / S) & 1;
                        if (!tc_wait_mbar(bar_empty + 8 * st, ph ^ 1, p, 2)) { ok = false; break; }
                        tc::mbar_arrive_expect_tx(bar_full + 8 * st, KLM_TC_CHUNK_BYTES);
                        tc::bulk_g2s(tc::smem_u32(ring + (size_t)st * KLM_TC_CHUNK_BYTES), src + (size_t)c * KLM_TC_CHUNK_BYTES,
                                     KLM_TC_CHUNK_BYTES, bar_full + 8 * st);
                        if (c0 == 0) tc_trace(p, tick, 1);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // =============================== MMA issuer ===============================
        if (lane == 0) {
            const uint32_t idesc = tc::idesc_f16(128, 2 * N);
            const uint32_t lbo_a = 128 * 16, lbo_b = 2 * N * 16;
            const uint32_t ring_a = tc::smem_u32(ring), w_a = tc::smem_u32(wsm);
            uint32_t it = 0;
            bool ok = true;
            for (int tick = 0; tick < nticks && ok; ++tick) {
                if (!tc_wait_mbar(bar_tempty, (tick & 1) ^ 1, p, 3)) break;
                tc::tc_fence_after();
                for (int c0 = 0; c0 < nchunks; ++c0, ++it) {
                    const int c = (c0 + (int)blockIdx.x) % nchunks;
                    const uint32_t st = it % S, ph = (it / S) & 1;
                    if (!tc_wait_mbar(bar_full + 8 * st, ph, p, 4)) { ok = false; break; }
                    tc_trace(p, tick, c0 == 0 ? 2 : 3);
                    tc::tc_fence_after();
#pragma unroll
                    for (int kk = 0; kk < KLM_TC_KC / 16; ++kk) {
                        const uint64_t ad = tc::smem_desc(ring_a + st * KLM_TC_CHUNK_BYTES + kk * 2 * lbo_a, lbo_a, 128);
                        const uint64_t bd = tc::smem_desc(w_a + (uint32_t)(c * (KLM_TC_KC / 8) + kk * 2) * lbo_b, lbo_b, 128);
                        tc::mma_f16(tmem_base, ad, bd, idesc, (c0 | kk) != 0 ? 1u : 0u);
                    }
                    tc::mma_commit(bar_empty + 8 * st);       // ring slot free once these MMAs have read it
                }
                if (ok) tc::mma_commit(bar_tfull);            // accumulators of this tick complete
            }
        }
    } else {
        // =============================== epilogue ===============================
        const uint32_t trow = tmem_base + ((uint32_t)(32 * q) << 16);
        float4 bias4[KLM_TC_OWN][2];
        if (role.kind == KLM_TC_IN) {
#pragma unroll
            for (int i = 0; i < KLM_TC_OWN; ++i)
                if (i < nown) {
                    const int u = 2 * (role.first_pair + own_first + i);
                    bias4[i][0] = __ldg(p.bias[l] + u); bias4[i][1] = __ldg(p.bias[l] + u + 1);
                }
        }
        uint32_t lit = 0;                                 // ring position of the load/store ingest
        for (int tick = 0; tick < nticks; ++tick) {
            // the producer thread polls the grid barrier; its mbarrier arrive (release.cta) hands the
            // acquired view of the other CTAs' writes (Z, logits) on to the epilogue threads
            int ok = tc_wait_mbar(bar_go, tick & 1, p, 5) ? 1 : 0;
            ok = __all_sync(0xffffffffu, ok);
            if (!ok) break;
            const int t = tick - role.sigma;
            const bool active = s < p.B && t >= 0 && t < len;
            const int par_in = tick & 1, par_out = (tick + 1) & 1;

            // ---- addends, fetched while the MMAs run ----
            float4 add4[KLM_TC_OWN][2];
#pragma unroll
            for (int i = 0; i < KLM_TC_OWN; ++i) add4[i][0] = add4[i][1] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (is_rec && active) {
                if (l == 0) {
                    // fused char + context embedding gather / concat / input contraction (rating.py:103-115,125)
                    int x = __ldg(p.chars + o0 + t);
                    x = min(max(x, 0), p.V - 1);
                    const bool real = p.n_real ? (t < __ldg(p.n_real + s)) : true;
                    int z[KLM_MAX_CTX];
                    for (int k = 0; k < p.Nc; ++k) {
                        z[k] = real ? __ldg(p.ctx + s * p.Nc + k) : 0;       // padding: context 0 (:1141-1143)
                        z[k] = min(max(z[k], 0), p.ctx_rows[k] - 1);
                    }
#pragma unroll
                    for (int i = 0; i < KLM_TC_OWN; ++i)
                        if (i < nown) {
                            const int u = 2 * (role.first_pair + own_first + i);
#pragma unroll
                            for (int e = 0; e < 2; ++e) {
                                float4 a = __ldg(p.P + (size_t)x * W + u + e);
                                for (int k = 0; k < p.Nc; ++k) {
                                    const float4 b = __ldg(p.Q[k] + (size_t)z[k] * W + u + e);
                                    a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
                                }
                                add4[i][e] = a;
                            }
                        }
                } else {
                    const float4 *zrow = reinterpret_cast<const float4 *>(p.Z + ((size_t)(l * 2 + par_in) * 64 + s) * 4 * W);
#pragma unroll
                    for (int i = 0; i < KLM_TC_OWN; ++i)
                        if (i < nown) {
                            const int u = 2 * (role.first_pair + own_first + i);
                            add4[i][0] = __ldcg(zrow + u); add4[i][1] = __ldcg(zrow + u + 1);
                        }
                }
            }

            // ---- A operand of this tick: the hidden vector, global (L2) -> ring, by the epilogue warps ----
            if (!p.ingest_bulk) {
                const uint4 *src = reinterpret_cast<const uint4 *>(p.hx + (size_t)(role.src_layer * 2 + par_in) * W * 128);
                const int et = tid - 64;                                   // 0..255; a chunk is 4 x 256 uint4
                uint4 ba[4], bb[4], bc[4];
                bool okl = true;
#define TC_LOAD(c0_, buf_)                                                                      \
                {                                                                               \
                    const uint4 *sp_ = src + (size_t)(((c0_) + (int)blockIdx.x) % nchunks) * 1024 + et; \
                    _Pragma("unroll") for (int j = 0; j < 4; ++j) buf_[j] = __ldcg(sp_ + 256 * j); \
                }
#define TC_STORE(c0_, buf_)                                                                     \
                {                                                                               \
                    const uint32_t st_ = lit % S, ph_ = (lit / S) & 1;                          \
                    okl = okl && tc_wait_mbar(bar_empty + 8 * st_, ph_ ^ 1, p, 2);              \
                    uint4 *dp_ = reinterpret_cast<uint4 *>(ring + (size_t)st_ * KLM_TC_CHUNK_BYTES) + et; \
                    _Pragma("unroll") for (int j = 0; j < 4; ++j) dp_[256 * j] = buf_[j];       \
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");                \
                    tc::mbar_arrive(bar_full + 8 * st_);                                        \
                    if ((c0_) == 0 && tid == 64) tc_trace(p, tick, 1);                          \
                    ++lit;                                                                      \
                }
                TC_LOAD(0, ba);
                if (nchunks > 1) TC_LOAD(1, bb);
                for (int c0 = 0; c0 < nchunks; c0 += 3) {
                    if (c0 + 2 < nchunks) TC_LOAD(c0 + 2, bc);
                    TC_STORE(c0, ba);
                    if (c0 + 3 < nchunks) TC_LOAD(c0 + 3, ba);
                    if (c0 + 1 < nchunks) TC_STORE(c0 + 1, bb);
                    if (c0 + 4 < nchunks) TC_LOAD(c0 + 4, bb);
                    if (c0 + 2 < nchunks) TC_STORE(c0 + 2, bc);
                }
#undef TC_LOAD
#undef TC_STORE
                if (!__all_sync(0xffffffffu, okl ? 1 : 0)) break;
            }

            // ---- accumulators ----
            ok = tc_wait_mbar(bar_tfull, tick & 1, p, 6) ? 1 : 0;
            ok = __all_sync(0xffffffffu, ok);
            if (!ok) break;
            if (tid == 64) tc_trace(p, tick, 4);
            tc::tc_fence_after();
            float acc[KLM_TC_OWN][8];
#pragma unroll
            for (int i = 0; i < KLM_TC_OWN; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
            // own pairs: this row's part (hi rows: hi.hi + hi.lo columns; lo rows: lo.hi columns)
#pragma unroll
            for (int i = 0; i < KLM_TC_OWN; ++i)
                if (i < nown) {
                    const int pr = own_first + i;
                    tc::tmem_ld8(trow + 8 * pr, acc[i]);
                    if (half == 0) {
                        float b[8];
                        tc::tmem_ld8(trow + N + 8 * pr, b);
                        tc::tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < 8; ++j) acc[i][j] += b[j];
                    } else {
                        tc::tmem_ld_wait();
                    }
                }
            // the other half's pairs: hand this row's part over through shared memory (both directions at once)
            {
                float *xs = xchg + (half ? 0 : KLM_TC_MAX_PAIRS / 2 * 8 * 64);
#pragma unroll
                for (int i = 0; i < KLM_TC_OWN; ++i)
                    if (i < nsend) {
                        const int pr = send_first + i, xi = pr - oth_first;
                        float v[8];
                        tc::tmem_ld8(trow + 8 * pr, v);
                        if (half == 0) {
                            float b[8];
                            tc::tmem_ld8(trow + N + 8 * pr, b);
                            tc::tmem_ld_wait();
#pragma unroll
                            for (int j = 0; j < 8; ++j) v[j] += b[j];
                        } else {
                            tc::tmem_ld_wait();
                        }
#pragma unroll
                        for (int j = 0; j < 8; ++j) xs[(xi * 8 + j) * 64 + s] = v[j];
                    }
            }
            tc::tc_fence_before();
            tc::mbar_arrive(bar_tempty);                      // this thread is done reading TMEM
            tc_bar_epi();
            {
                const float *xr = xchg + (half ? KLM_TC_MAX_PAIRS / 2 * 8 * 64 : 0);
#pragma unroll
                for (int i = 0; i < KLM_TC_OWN; ++i)
                    if (i < nown) {
                        const int xi = own_first + i - my_first;
#pragma unroll
                        for (int j = 0; j < 8; ++j) acc[i][j] += xr[(xi * 8 + j) * 64 + s];
                    }
                // the hand-over buffer aliases the ring: order these generic accesses before the next bulk copies
                if (p.ingest_bulk) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            }

            // ---- finish and hand over ----
            if (tid == 64) tc_trace(p, tick, 5);
            if (is_rec) {
                // LSTM cell for all owned units, staged so that the MUFU chains of different units overlap
                float e1[KLM_TC_OWN][2], e2[KLM_TC_OWN][2], e3[KLM_TC_OWN][2], e4[KLM_TC_OWN][2];
#pragma unroll
                for (int i = 0; i < KLM_TC_OWN; ++i)
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const float4 a = add4[i][e];
                        e1[i][e] = tc_exp(-tc_clamp(acc[i][4 * e] + a.x));
                        e2[i][e] = tc_exp(-tc_clamp(acc[i][4 * e + 1] + a.y));
                        e3[i][e] = tc_exp(-2.f * tc_clamp(acc[i][4 * e + 2] + a.z));
                        e4[i][e] = tc_exp(-tc_clamp(acc[i][4 * e + 3] + a.w));
                    }
#pragma unroll
                for (int i = 0; i < KLM_TC_OWN; ++i)
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const float ig = __fdividef(1.f - e3[i][e], (1.f + e1[i][e]) * (1.f + e3[i][e]));
                        const float f = __fdividef(1.f, 1.f + e2[i][e]);
                        e1[i][e] = fmaf(f, c_reg[i][e], ig);                       // new cell state
                        e2[i][e] = tc_exp(-2.f * tc_clamp(e1[i][e]));
                    }
#pragma unroll
                for (int i = 0; i < KLM_TC_OWN; ++i)
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const float hn = __fdividef(1.f - e2[i][e], (1.f + e4[i][e]) * (1.f + e2[i][e]));
                        if (active) { c_reg[i][e] = e1[i][e]; h_reg[i][e] = hn; }
                    }
#pragma unroll
                for (int i = 0; i < KLM_TC_OWN; ++i)
                    if (i < nown) {
                        const int u = 2 * (role.first_pair + own_first + i);
                        __half hi0, lo0, hi1, lo1;
                        tc_split(h_reg[i][0], hi0, lo0); tc_split(h_reg[i][1], hi1, lo1);
                        __half *dst = p.hx + ((size_t)(l * 2 + par_out) * (W / 8) + u / 8) * (128 * 8) + (u & 7);
                        *reinterpret_cast<__half2 *>(dst + s * 8) = __halves2half2(hi0, hi1);
                        *reinterpret_cast<__half2 *>(dst + (64 + s) * 8) = __halves2half2(lo0, lo1);
                    }
                if (p.ingest_bulk) tc::fence_proxy_async();
            } else if (role.kind == KLM_TC_IN) {
                float4 *zrow = reinterpret_cast<float4 *>(p.Z + ((size_t)(l * 2 + par_out) * 64 + s) * 4 * W);
#pragma unroll
                for (int i = 0; i < KLM_TC_OWN; ++i)
                    if (i < nown) {
                        const int u = 2 * (role.first_pair + own_first + i);
#pragma unroll
                        for (int e = 0; e < 2; ++e) {
                            const float4 b = bias4[i][e];
                            __stcg(zrow + u + e, make_float4(acc[i][4 * e] + b.x, acc[i][4 * e + 1] + b.y,
                                                             acc[i][4 * e + 2] + b.z, acc[i][4 * e + 3] + b.w));
                        }
                    }
            } else {   // PROJ: logits of vocabulary rows 8*(first_pair + pair) ..
                float4 *lrow = reinterpret_cast<float4 *>(p.logits + ((size_t)par_out * 64 + s) * p.Vp);
#pragma unroll
                for (int i = 0; i < KLM_TC_OWN; ++i)
                    if (i < nown) {
                        const int g8 = role.first_pair + own_first + i;
                        __stcg(lrow + 2 * g8, make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]));
                        __stcg(lrow + 2 * g8 + 1, make_float4(acc[i][4], acc[i][5], acc[i][6], acc[i][7]));
                    }
            }
            if (tid == 64) tc_trace(p, tick, 6);
            tc_bar_epi();
            if (tid == 64) { tc::red_release_add(p.barrier, 1u); tc_trace(p, tick, 7); }
        }
        // ---- write the final state back ----
        if (is_rec && s < p.B) {
            float *st = p.state + ((size_t)p.slots_out[s] * p.D + l) * 2 * W;
#pragma unroll
            for (int i = 0; i < KLM_TC_OWN; ++i)
                if (i < nown) {
                    const int u = 2 * (role.first_pair + own_first + i);
                    st[u] = h_reg[i][0]; st[u + 1] = h_reg[i][1];
                    st[W + u] = c_reg[i][0]; st[W + u + 1] = c_reg[i][1];
                }
        }
    }

    tc::tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        __syncwarp();
        tc::tmem_dealloc(tmem_base, KLM_TC_TMEM_COLS);
    }
}

cudaError_t klm_launch_seq_tc(const KlmTcParams &p, int grid, size_t smem, cudaStream_t stream) {
    void *args[] = { const_cast<KlmTcParams *>(&p) };
    cudaError_t e = cudaFuncSetAttribute((const void *)klm_seq_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    return cudaLaunchCooperativeKernel((const void *)klm_seq_tc, dim3(grid), dim3(KLM_TC_THREADS), args, smem, stream);
}

// ---------------------------------------------------------------------------------------------
// Self-test of the tcgen05 building blocks: D[128 x N] = A[128 x K] . B[N x K]^T for fp16 operand
// images in the layout above (A through bulk async copies, B through generic stores).
__global__ void __launch_bounds__(128, 1) klm_umma_selftest(const __half *A, const __half *B, float *Dout, int N, int K,
                                                             int *err) {
    extern __shared__ __align__(1024) unsigned char smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t a_bytes = (uint32_t)K * 128 * 2, b_bytes = (uint32_t)K * N * 2;
    unsigned char *sa = smem, *sb = smem + a_bytes;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + a_bytes + b_bytes);
    const uint32_t bar_full = tc::smem_u32(bars), bar_done = tc::smem_u32(bars + 1);
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2);
    if (warp == 0) { tc::tmem_alloc(tc::smem_u32(tmem_slot), 256); tc::tmem_relinquish(); }
    if (tid == 32) { tc::mbar_init(bar_full, 1); tc::mbar_init(bar_done, 1); tc::fence_mbar_init(); }
    for (uint32_t i = tid; i < b_bytes / 16; i += 128) reinterpret_cast<uint4 *>(sb)[i] = reinterpret_cast<const uint4 *>(B)[i];
    tc::fence_proxy_async();
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (tid == 0) {
        tc::mbar_arrive_expect_tx(bar_full, a_bytes);
        for (uint32_t off = 0; off < a_bytes; off += 16384)
            tc::bulk_g2s(tc::smem_u32(sa + off), reinterpret_cast<const unsigned char *>(A) + off, min(16384u, a_bytes - off), bar_full);
    }
    if (tid == 32) {
        unsigned n = 0;
        while (!tc::mbar_try_wait(bar_full, 0)) if (++n > (1u << 22)) { atomicExch(err, 1); break; }
        tc::tc_fence_after();
        const uint32_t idesc = tc::idesc_f16(128, N);
        const uint32_t lbo_a = 128 * 16, lbo_b = N * 16;
        for (int k = 0; k < K / 16; ++k)
            tc::mma_f16(tmem_base, tc::smem_desc(tc::smem_u32(sa) + k * 2 * lbo_a, lbo_a, 128),
                        tc::smem_desc(tc::smem_u32(sb) + k * 2 * lbo_b, lbo_b, 128), idesc, k != 0 ? 1u : 0u);
        tc::mma_commit(bar_done);
    }
    __syncwarp();
    {
        unsigned n = 0;
        while (!tc::mbar_try_wait(bar_done, 0)) if (++n > (1u << 22)) { atomicExch(err, 2); break; }
    }
    __syncwarp();
    tc::tc_fence_after();
    const uint32_t trow = tmem_base + ((uint32_t)(32 * warp) << 16);
    for (int c = 0; c < N; c += 8) {
        float v[8];
        tc::tmem_ld8(trow + c, v);
        tc::tmem_ld_wait();
        for (int j = 0; j < 8; ++j) Dout[(size_t)(32 * warp + lane) * N + c + j] = v[j];
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 0) { __syncwarp(); tc::tmem_dealloc(tmem_base, 256); }
}

cudaError_t klm_launch_umma_selftest(const __half *A, const __half *B, float *D, int N, int K, int *err, cudaStream_t stream) {
    const size_t smem = (size_t)K * 128 * 2 + (size_t)K * N * 2 + 64;
    cudaError_t e = cudaFuncSetAttribute((const void *)klm_umma_selftest, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    klm_umma_selftest<<<1, 128, smem, stream>>>(A, B, D, N, K, err);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// Self-test of the CTA-pair MMA (cta_group::2, M = 128 over two CTAs): CTA r holds rows 64r..64r+63 of A
// ([K/8][64][8] halves) and rows (N/2)r.. of B ([K/8][N/2][8]); every CTA dumps its TMEM (128 lanes x N/2
// columns) so that the host can check the accumulator layout.
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
klm_umma2_selftest(const __half *A, const __half *B, float *Dout, int N, int K, int *err) {
    extern __shared__ __align__(1024) unsigned char smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t rank = tc::cluster_ctarank();
    const uint32_t a_bytes = (uint32_t)K * 64 * 2, b_bytes = (uint32_t)K * (N / 2) * 2;
    unsigned char *sa = smem, *sb = smem + a_bytes;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + a_bytes + b_bytes);
    const uint32_t bar_done = tc::smem_u32(bars);
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2);
    if (warp == 0) { tc::tmem_alloc2(tc::smem_u32(tmem_slot), 256); tc::tmem_relinquish2(); }
    if (tid == 32) { tc::mbar_init(bar_done, 1); tc::fence_mbar_init(); }
    const uint4 *ga = reinterpret_cast<const uint4 *>(reinterpret_cast<const unsigned char *>(A) + (size_t)rank * a_bytes);
    const uint4 *gb = reinterpret_cast<const uint4 *>(reinterpret_cast<const unsigned char *>(B) + (size_t)rank * b_bytes);
    for (uint32_t i = tid; i < a_bytes / 16; i += 128) reinterpret_cast<uint4 *>(sa)[i] = ga[i];
    for (uint32_t i = tid; i < b_bytes / 16; i += 128) reinterpret_cast<uint4 *>(sb)[i] = gb[i];
    tc::fence_proxy_async();
    tc::tc_fence_before();
    __syncthreads();
    tc::cluster_sync();                                  // both CTAs: operands in place, TMEM allocated, barriers ready
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (rank == 0 && tid == 32) {
        const uint32_t idesc = tc::idesc_f16(128, N);
        const uint32_t lbo_a = 64 * 16, lbo_b = (N / 2) * 16;
        for (int k = 0; k < K / 16; ++k)
            tc::mma2_f16(tmem_base, tc::smem_desc(tc::smem_u32(sa) + k * 2 * lbo_a, lbo_a, 128),
                         tc::smem_desc(tc::smem_u32(sb) + k * 2 * lbo_b, lbo_b, 128), idesc, k != 0 ? 1u : 0u);
        tc::mma2_commit(bar_done, 3);
    }
    __syncwarp();
    {
        unsigned n = 0;
        while (!tc::mbar_try_wait(bar_done, 0)) if (++n > (1u << 22)) { atomicExch(err, 2 + (int)rank); break; }
    }
    __syncwarp();
    tc::tc_fence_after();
    const uint32_t trow = tmem_base + ((uint32_t)(32 * warp) << 16);
    for (int c = 0; c < N / 2; c += 8) {
        float v[8];
        tc::tmem_ld8(trow + c, v);
        tc::tmem_ld_wait();
        for (int j = 0; j < 8; ++j) Dout[((size_t)rank * 128 + 32 * warp + lane) * (N / 2) + c + j] = v[j];
    }
    tc::tc_fence_before();
    __syncthreads();
    tc::cluster_sync();
    if (warp == 0) { __syncwarp(); tc::tmem_dealloc2(tmem_base, 256); }
}

cudaError_t klm_launch_umma2_selftest(const __half *A, const __half *B, float *D, int N, int K, int *err, cudaStream_t stream) {
    const size_t smem = (size_t)K * 64 * 2 + (size_t)K * (N / 2) * 2 + 64;
    cudaError_t e = cudaFuncSetAttribute((const void *)klm_umma2_selftest, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    klm_umma2_selftest<<<2, 128, smem, stream>>>(A, B, D, N, K, err);
    return cudaGetLastError();
}
