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
// Launches of >= 64 positions run with skew 3 instead (KlmTcParams::skew): REC l at 3l, IN l at 3l-2,
// PROJ at 3D-2, LSE at 3D; IN / PROJ CTAs signal the barrier as soon as their MMAs have read the operand,
// their epilogue runs into the next tick, Z and the logits live in 4 slots and are consumed two ticks
// after the GEMM -- the recurrence (REC) never waits for an IN / PROJ epilogue.
//
// Warp roles in a GEMM CTA (320 threads): warp 0 lane 0 = producer (grid barrier wait + bulk
// copies), warp 1 lane 0 = MMA issuer, warps 2..9 = epilogue (TMEM -> registers, hi/lo halves
// exchanged through shared memory, LSTM cell in registers, hand-over stores).
//
// Publication of the hidden vector (role.pub): the readers' bulk copies run at ~46 B/clk per SM when
// the lines were last written by full-sector stores but only ~28 B/clk (19 in this kernel) after
// scattered 4-byte stores (scripts/ubench/fetch_paths.cu), so REC CTAs own slices aligned to 8 units
// (one 16-byte row piece of the operand layout), stage the new hi/lo halves in shared memory and write
// them out as whole 2 KB blocks (coalesced 16-byte stores or one bulk async store).
#include <cstdio>
#include "klm_tc_dev.cuh"

// ---------------------------------------------------------------------------------------------
// Producer and MMA-issuer loops for compile-time ring geometries (KC k per chunk, NCH chunks per tick, S slots).
// Both roles are ONE thread each, and a lone thread retires a dependent instruction only every ~5 clk; with
// run-time ring arithmetic the per-chunk bookkeeping (~400 clk, measured with the MMAs and copies switched off,
// KLM_TC_DEBUG_SKIP) cost more than the four MMAs of the chunk (~290 clk), and because the MMA queue is shallow
// the two add up.  Here slot numbers, shared-memory offsets and the K loop are unrolled constants; the only
// running state is the chunk rotation and one phase bit per slot.
//
// STREAM = streamed-weights form: a ring slot is [A chunk | the same k-rows of this CTA's B image]; both copies
// complete on the slot's barrier.  The B parts of the first ring round do not depend on the tick, so they are
// issued before the grid barrier and land while the previous tick's epilogue and hand-over are still going on.
// With more than 128 columns per half the chunk is multiplied in two MMAs (hi columns, lo columns) of N = 8 * pairs.
struct TcLoopCtx {
    uint32_t ring_a, bar_full, bar_empty, bar_tfull, bar_tempty, bar_go, tmem_base, wsm_a, bar_epid;
    int nticks, N, crot, src_layer, W, guard;
    unsigned G;
    uint32_t slot_bytes, cbb;              // STREAM: bytes per ring slot (A chunk + B chunk), bytes of a B chunk
    const unsigned char *bsrc;             // STREAM: this CTA's B image in global memory
};

template <int KC, int NCH, int S, bool TRACE, bool STREAM>
__device__ __noinline__ void tc_produce_fixed(const KlmTcParams &p, const TcLoopCtx &x) {
    constexpr uint32_t CB = KC * 256;
    // everything the loop needs sits in registers: the asm statements clobber memory, so anything left in the
    // context struct or the kernel parameters would be re-loaded after each of them
    const uint32_t ring_a = x.ring_a, bar_full = x.bar_full, bar_empty = x.bar_empty, bar_go = x.bar_go;
    const uint32_t slot_bytes = STREAM ? x.slot_bytes : CB, cbb = STREAM ? x.cbb : 0u;
    const unsigned char *bsrc = x.bsrc;
    const int nticks = x.nticks, crot = x.crot;
    const unsigned G = x.G;
    const int guard = x.guard;
    const uint32_t bar_epid = x.bar_epid;
    const int trace_mode = p.trace_mode, skip = TRACE ? p.debug_skip : 0, cfence = p.consumer_fence;   // (timing experiments: traced instantiations only)
    const unsigned char *hx0 = reinterpret_cast<const unsigned char *>(p.hx + (size_t)(x.src_layer * 2) * x.W * 128);
    const size_t hx_par = (size_t)x.W * 128 * sizeof(__half);
    uint32_t phbits = 0;                                   // bit st: uses of slot st so far, mod 2
    const bool lane0 = (threadIdx.x & 31) == 0;
    constexpr int NPRE = S < NCH ? S : NCH;
    for (int tick = 0; tick < nticks; ++tick) {
        // the whole warp runs this loop converged; lane 0 polls the grid barrier and issues.
        // The first ring round is armed before the grid barrier (its slots are free since the previous tick's
        // MMAs), so that after the barrier the copies go out back to back.
#pragma unroll
        for (int c0 = 0; c0 < NPRE; ++c0) {
            int okw = tc_wait_mbar(bar_empty + 8 * c0, ((phbits >> c0) & 1u) ^ 1u, p, 2) ? 1 : 0;
            if (!__all_sync(0xffffffffu, okw)) return;
            if (tc::elect_one() && !(skip & 2)) {
                tc::mbar_arrive_expect_tx(bar_full + 8 * c0, CB + cbb);
                if (STREAM) {
                    const int cb = crot + c0 >= NCH ? crot + c0 - NCH : crot + c0;
                    tc::bulk_g2s(ring_a + c0 * slot_bytes + CB, bsrc + (size_t)cb * cbb, cbb, bar_full + 8 * c0);
                }
            }
            __syncwarp();
        }
        if (guard && tick > 0) {       // (IN / PROJ with the early signal: the previous epilogue has left the aliased buffers)
            int oke = tc_wait_mbar(bar_epid, (uint32_t)(tick - 1) & 1u, p, 7) ? 1 : 0;
            if (!__all_sync(0xffffffffu, oke)) return;
        }
        int okg = 1;
        if (lane0) {
            okg = tc_wait_grid(p, (unsigned)(tick + 1) * G, 1) ? 1 : 0;
            if (okg) {
                if (TRACE && trace_mode == 3) tc_trace(p, tick, 0);
                tc::mbar_arrive(bar_go);
                if (TRACE && trace_mode == 0) tc_trace(p, tick, 0);
                if (TRACE && trace_mode == 3) tc_trace(p, tick, 1);
                if (cfence) tc::fence_proxy_async();
            }
        }
        if (!__shfl_sync(0xffffffffu, okg, 0)) return;
        if (TRACE && trace_mode == 3 && lane0) tc_trace(p, tick, 2);
        const unsigned char *src = hx0 + ((tick & 1) ? hx_par : 0);
        int c = crot;
#pragma unroll
        for (int c0 = 0; c0 < NCH; ++c0) {
            const int st = c0 % S;
            if (c0 >= NPRE) {
                int okw = tc_wait_mbar(bar_empty + 8 * st, ((phbits >> st) & 1u) ^ 1u, p, 2) ? 1 : 0;
                if (!__all_sync(0xffffffffu, okw)) return;
            }
            if (tc::elect_one()) {
                if (TRACE && trace_mode == 3 && c0 == 0) tc_trace(p, tick, 3);
                if (skip & 2) tc::mbar_arrive(bar_full + 8 * st);
                else {
                    if (c0 >= NPRE) tc::mbar_arrive_expect_tx(bar_full + 8 * st, CB + cbb);
                    tc::bulk_g2s(ring_a + st * slot_bytes, src + (size_t)c * CB, CB, bar_full + 8 * st);
                    if (STREAM && c0 >= NPRE)
                        tc::bulk_g2s(ring_a + st * slot_bytes + CB, bsrc + (size_t)c * cbb, cbb, bar_full + 8 * st);
                }
                if (TRACE && trace_mode == 3 && c0 < 4) tc_trace(p, tick, 4 + c0);
                if (TRACE) {
                    if (trace_mode == 2) tc_trace(p, tick, c0 & 7);
                    else if (c0 == 0 && trace_mode == 0) tc_trace(p, tick, 1);
                }
            }
            __syncwarp();
            phbits ^= 1u << st;
            c = (c + 1 == NCH) ? 0 : c + 1;
        }
    }
}

template <int KC, int NCH, int S, bool TRACE, bool STREAM>
__device__ __noinline__ void tc_mma_fixed(const KlmTcParams &p, const TcLoopCtx &x) {
    constexpr uint32_t CB = KC * 256;
    const bool split = STREAM && 2 * x.N > 256;            // two MMAs per k step: hi columns, lo columns
    const uint32_t idesc = tc::idesc_f16(128, split ? x.N : 2 * x.N);
    const uint32_t lbo_a = 128 * 16, lbo_b = 2 * x.N * 16;
    const uint32_t slot16 = (STREAM ? x.slot_bytes : CB) >> 4;
    const uint64_t a_base = tc::smem_desc(x.ring_a, lbo_a, 128);
    const uint64_t b_base = tc::smem_desc(STREAM ? x.ring_a + CB : x.wsm_a, lbo_b, 128);
    constexpr uint32_t a_k16 = (2 * 128 * 16) >> 4;
    const uint32_t b_k16 = (2 * lbo_b) >> 4, b_chunk = ((uint32_t)(KC / 8) * lbo_b) >> 4;
    const uint32_t b_lo = (uint32_t)x.N;                   // lo rows start N * 16 bytes into a k group; also the TMEM column
    const uint32_t bar_full = x.bar_full, bar_empty = x.bar_empty, bar_tfull = x.bar_tfull, bar_tempty = x.bar_tempty;
    const uint32_t tmem_base = x.tmem_base;
    const int nticks = x.nticks, crot = x.crot;
    const int trace_mode = p.trace_mode, skip = TRACE ? p.debug_skip : 0;
    uint32_t phbits = 0;
    for (int tick = 0; tick < nticks; ++tick) {
        // the whole warp runs this loop converged; the elected lane (always lane 0) issues
        int okw = tc_wait_mbar(bar_tempty, (tick & 1) ^ 1, p, 3) ? 1 : 0;
        if (!__all_sync(0xffffffffu, okw)) return;
        tc::tc_fence_after();
        int c = crot;
#pragma unroll
        for (int c0 = 0; c0 < NCH; ++c0) {
            const int st = c0 % S;
            okw = tc_wait_mbar(bar_full + 8 * st, (phbits >> st) & 1u, p, 4) ? 1 : 0;
            if (!__all_sync(0xffffffffu, okw)) return;
            const uint64_t ad = a_base + (uint32_t)(st * slot16);
            const uint64_t bd = STREAM ? b_base + (uint32_t)(st * slot16) : b_base + (uint32_t)c * b_chunk;
            if (tc::elect_one()) {
                if (TRACE) {
                    if (trace_mode == 1) tc_trace(p, tick, c0 & 7);
                    else if (trace_mode == 0 && (c0 == 0 || c0 == NCH - 1)) tc_trace(p, tick, c0 == 0 ? 2 : 3);
                }
                if (!(skip & 1)) {
                    if (STREAM && split) {
#pragma unroll
                        for (int kk = 0; kk < KC / 16; ++kk) {
                            tc::mma_f16(tmem_base, ad + kk * a_k16, bd + kk * b_k16, idesc, (c0 | kk) != 0 ? 1u : 0u);
                            tc::mma_f16(tmem_base + b_lo, ad + kk * a_k16, bd + kk * b_k16 + b_lo, idesc, (c0 | kk) != 0 ? 1u : 0u);
                        }
                    } else {
#pragma unroll
                        for (int kk = 0; kk < KC / 16; ++kk)
                            tc::mma_f16(tmem_base, ad + kk * a_k16, bd + kk * b_k16, idesc, (c0 | kk) != 0 ? 1u : 0u);
                    }
                }
                tc::mma_commit(bar_empty + 8 * st);       // ring slot free once these MMAs have read it
                if (c0 == NCH - 1) tc::mma_commit(bar_tfull);   // accumulators of this tick complete
            }
            __syncwarp();
            phbits ^= 1u << st;
            c = (c + 1 == NCH) ? 0 : c + 1;
        }
    }
}

#define TC_FIXED_GEOMS(F) F(64, 8, 4) F(64, 8, 3) F(32, 16, 4) F(64, 4, 4) F(128, 4, 3) F(128, 4, 2)
#define TC_STREAM_GEOMS(F) F(64, 8, 3) F(64, 12, 3) F(64, 16, 3) F(64, 24, 3) F(64, 32, 3)     // widths 512, 768, 1024, 1536, 2048

template <bool STREAM, bool TRACE>
static __device__ __forceinline__ bool tc_run_fixed(bool producer, const KlmTcParams &p, const TcLoopCtx &x, int kc, int nch, int S) {
#define TC_TRY(KC_, NCH_, S_)                                                                                   \
    if (kc == KC_ && nch == NCH_ && S == S_) {                                                                  \
        if (producer) tc_produce_fixed<KC_, NCH_, S_, TRACE, STREAM>(p, x);                                     \
        else tc_mma_fixed<KC_, NCH_, S_, TRACE, STREAM>(p, x);                                                  \
        return true;                                                                                            \
    }
    if (STREAM) { TC_STREAM_GEOMS(TC_TRY) }
    else { TC_FIXED_GEOMS(TC_TRY) }
#undef TC_TRY
    return false;
}

// One epilogue round = up to RP = 4 * OWN pairs of the CTA's slice.  The pairs of a round are split between the hi-row
// and the lo-row threads, and between the two warps that share a TMEM lane quarter; all indices below are relative
// to the round's first pair `pb`.  (Recomputed where it is needed: an array of these per thread cost 7 registers a round.)
struct TcRound { int pb, my_first, oth_first, own_first, nown, send_first, nsend; };
template <int RP>
static __device__ __forceinline__ TcRound tc_round(int NP, int r, int half, int ch) {
    TcRound o;
    o.pb = r * RP;
    const int npr = min(max(NP - o.pb, 0), RP);
    const int Ph = (npr + 1) / 2;                    // pairs [0,Ph) are finished by hi-row threads, the rest by lo-row threads
    const int my_n = half ? npr - Ph : Ph, oth_n = half ? Ph : npr - Ph;
    o.my_first = half ? Ph : 0;
    o.oth_first = half ? 0 : Ph;
    o.own_first = o.my_first + (ch ? (my_n + 1) / 2 : 0);
    o.nown = ch ? my_n - (my_n + 1) / 2 : (my_n + 1) / 2;
    o.send_first = o.oth_first + (ch ? (oth_n + 1) / 2 : 0);
    o.nsend = ch ? oth_n - (oth_n + 1) / 2 : (oth_n + 1) / 2;
    return o;
}

// ---------------------------------------------------------------------------------------------
// NR = epilogue rounds (pairs per CTA <= NR * 4 * OWN), STREAM = streamed-weights form.
// TRACE = timeline instrumentation compiled in (klm_tc_trace armed): the production instantiations carry none of it --
// a handful of run-time-disabled checks in the epilogue cost the resident kernel 3 % on cfg2.
// OWN = pairs of units one epilogue thread finishes per round (2 covers slices of <= 8 pairs), REC = the CTA is a
// recurrence stage.  Compile-time so that a REC CTA carries no IN / PROJ state and vice versa: the kernel runs at its
// register limit (10 warps: 168 registers), and a spilled value costs an L2 round trip in the tick loop -- the acquire
// polls of the grid barrier invalidate L1 every time (12 spilled floats were 5 % on cfg2).
template <int NR, bool STREAM, bool TRACE, int OWN, bool REC>
static __device__ __forceinline__ void tc_seq_role(const KlmTcParams &p, const KlmTcRole &role, unsigned char *smem) {

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int W = p.W, S = role.stages, kc = role.kc;
    const unsigned G = gridDim.x;
    const int nticks = p.T_max + p.skew * p.D;
    const int NP = role.npairs, N = 8 * NP;
    const int nchunks = W / kc;
    const uint32_t chunk_bytes = (uint32_t)kc * 256u;
    constexpr int TMEM_COLS = STREAM ? 512 : KLM_TC_TMEM_COLS;

    // ---- shared memory carve-up ----
    // resident form: [ring, aliased by hand-over + publication staging][barriers][B image]
    // streamed form: [hand-over][publication staging][ring of (A chunk | B chunk) slots][barriers]
    unsigned char *ring = smem + (STREAM ? role.ring_ofs : 0);           // S chunks of the A operand
    float *xchg = reinterpret_cast<float *>(smem);                       // hi<->lo hand-over
    __half *pubbuf = reinterpret_cast<__half *>(smem + KLM_TC1_PUB_OFS); // publication staging
    float4 *zstage = reinterpret_cast<float4 *>(smem + KLM_TC1_PUB_OFS); // IN / PROJ with pub == 2: the slice of Z / of the logits
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + role.ring_bytes);
    unsigned char *wsm = smem + role.ring_bytes + 128;                   // B image (resident form)
    const uint32_t bar_full = tc::smem_u32(bars), bar_empty = tc::smem_u32(bars + 4);
    const uint32_t bar_tfull = tc::smem_u32(bars + 8), bar_tempty = tc::smem_u32(bars + 9);
    const uint32_t bar_go = tc::smem_u32(bars + 10);       // producer -> epilogue: grid barrier of this tick passed
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 11);
    const uint32_t bar_epid = tc::smem_u32(bars + 12);     // epilogue -> producer: done with the buffers that alias the ring
    if (warp == 0) {
        tc::tmem_alloc(tc::smem_u32(tmem_slot), TMEM_COLS);
        tc::tmem_relinquish();
    }
    if (tid == 32) {
        for (int i = 0; i < S; ++i) { tc::mbar_init(bar_full + 8 * i, 1); tc::mbar_init(bar_empty + 8 * i, 1); }
        tc::mbar_init(bar_tfull, 1);
        tc::mbar_init(bar_tempty, 256);
        tc::mbar_init(bar_go, 1);
        tc::mbar_init(bar_epid, 1);
        tc::fence_mbar_init();
    }
    if (!STREAM) {   // B image: global -> shared, once per launch
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
    constexpr int RP = 4 * OWN;                      // pairs per epilogue round
    const TcRound rd0 = tc_round<RP>(NP, 0, half, ch);
    auto rd = [&](const int r) { return r == 0 ? rd0 : tc_round<RP>(NP, r, half, ch); };      // (later rounds: recomputed)
    const bool is_epi = warp >= 2;
    constexpr bool is_rec = REC;
    const int l = role.layer;
    // wavefront stage of this CTA (see KlmTcParams::skew)
    const int sigma = is_rec ? p.skew * l : role.kind == KLM_TC_IN ? p.skew * l - (p.skew - 1) : p.skew * p.D - (p.skew - 1);
    // skew 3: an IN / PROJ CTA signals the grid barrier once its MMAs have read the operand (the only thing of this tick the
    // other CTAs wait for: the buffer it read is overwritten next tick); what it writes is consumed two ticks later and is
    // covered by the signal of the next tick.  The recurrence (REC) no longer waits for the IN / PROJ epilogues.
    const bool early = p.skew == 3 && !is_rec;
    // resident form: hand-over buffer and staging alias the ring, so with the early signal the producer must not refill
    // the ring before the epilogue of the previous tick is through
    const bool guard = early && !STREAM;
    // IN / PROJ: slice of Z / of the logits staged in shared memory and written by one bulk async store.  Pays when the
    // signal follows the stores (skew 2: cfg2 12.88 M -> 13.38 M chars/s); with the early signal the stores are off the
    // critical path and plain coalesced stores are as good or better (14.37 M against 14.23 M)
    const bool zpub = !is_rec && role.pub == 2 && (!early || p.zpub_early);

    float c_reg[NR][OWN][2], h_reg[NR][OWN][2];
    long long o0 = 0;
    int len = 0;
    if (is_epi) {
        if (s < p.B) { o0 = p.offsets[s]; len = (int)(p.offsets[s + 1] - o0); }
#pragma unroll
        for (int r = 0; r < NR; ++r)
#pragma unroll
            for (int i = 0; i < OWN; ++i) {
                c_reg[r][i][0] = c_reg[r][i][1] = h_reg[r][i][0] = h_reg[r][i][1] = 0.f;
                if (is_rec && i < rd(r).nown) {
                    const int u = 2 * (role.first_pair + rd(r).pb + rd(r).own_first + i);
                    if (s < p.B) {
                        const int slot = p.slots_in[s];
                        if (slot >= 0) {
                            const float *st = p.state + ((size_t)slot * p.D + l) * 2 * W;
                            h_reg[r][i][0] = st[u]; h_reg[r][i][1] = st[u + 1];
                            c_reg[r][i][0] = st[W + u]; c_reg[r][i][1] = st[W + u + 1];
                        }
                    }
                    // publish the carried hidden state for tick 0 (parity 0)
                    __half hi0, lo0, hi1, lo1;
                    tc_split(h_reg[r][i][0], hi0, lo0); tc_split(h_reg[r][i][1], hi1, lo1);
                    __half *dst = p.hx + ((size_t)(l * 2 + 0) * (W / 8) + u / 8) * (128 * 8) + (u & 7);
                    *reinterpret_cast<__half2 *>(dst + s * 8) = __halves2half2(hi0, hi1);
                    *reinterpret_cast<__half2 *>(dst + (64 + s) * 8) = __halves2half2(lo0, lo1);
                }
            }
        tc::fence_proxy_async();                    // generic global writes -> later bulk (async proxy) reads
    }
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (tid == 0) tc::red_release_add(p.barrier, 1u);
    TcLoopCtx lx;
    lx.ring_a = tc::smem_u32(ring); lx.bar_full = bar_full; lx.bar_empty = bar_empty; lx.bar_tfull = bar_tfull;
    lx.bar_tempty = bar_tempty; lx.bar_go = bar_go; lx.tmem_base = tmem_base; lx.wsm_a = tc::smem_u32(wsm);
    lx.nticks = nticks; lx.N = N; lx.crot = role.rot % nchunks; lx.src_layer = role.src_layer; lx.W = W;
    lx.G = G; lx.guard = guard ? 1 : 0; lx.bar_epid = bar_epid;
    lx.cbb = (uint32_t)kc * 32u * (uint32_t)NP;            // kc/8 k groups x 2N rows x 16 bytes
    lx.slot_bytes = chunk_bytes + lx.cbb;
    lx.bsrc = p.wimg + role.wofs;

    if (warp == 0) {
        // =============================== producer ===============================
        // (one thread: everything in these loops is a dependent chain of ~5 clk per instruction, so the ring
        //  position, phase and chunk index are running counters -- an integer division per chunk costs more
        //  than the copy itself)
        if (!tc_run_fixed<STREAM, TRACE>(true, p, lx, kc, nchunks, S) && lane == 0) {
            if (STREAM) { atomicExch(p.err, 90); return; }   // the streamed form has compile-time geometries only
            uint32_t st = 0, ph = 0, ring_ofs = 0;
            const int crot = role.rot % nchunks;                       // spread the CTAs over the chunks
            const uint32_t ring_a = tc::smem_u32(ring);
            bool ok = true;
            for (int tick = 0; tick < nticks && ok; ++tick) {
                if (guard && tick > 0 && !tc_wait_mbar(bar_epid, (uint32_t)(tick - 1) & 1u, p, 7)) break;
                if (!tc_wait_grid(p, (unsigned)(tick + 1) * G, 1)) break;
                tc::mbar_arrive(bar_go);
                if (TRACE && p.trace_mode == 0) tc_trace(p, tick, 0);
                // operand fetch by the bulk-copy engine
                if (p.consumer_fence) tc::fence_proxy_async();   // generic global writes of other CTAs -> async-proxy reads
                const unsigned char *src = reinterpret_cast<const unsigned char *>(
                    p.hx + (size_t)(role.src_layer * 2 + (tick & 1)) * W * 128);
                int c = crot;
                for (int c0 = 0; c0 < nchunks; ++c0) {
                    if (!tc_wait_mbar(bar_empty + 8 * st, ph ^ 1, p, 2)) { ok = false; break; }
                    if (p.debug_skip & 2) tc::mbar_arrive(bar_full + 8 * st);
                    else {
                        tc::mbar_arrive_expect_tx(bar_full + 8 * st, chunk_bytes);
                        tc::bulk_g2s(ring_a + ring_ofs, src + (size_t)c * chunk_bytes, chunk_bytes, bar_full + 8 * st);
                    }
                    if (TRACE && p.trace_mode == 2) tc_trace(p, tick, c0 & 7);
                    else if (TRACE && c0 == 0 && p.trace_mode == 0) tc_trace(p, tick, 1);
                    ring_ofs += chunk_bytes;
                    if (++st == (uint32_t)S) { st = 0; ph ^= 1; ring_ofs = 0; }
                    if (++c == nchunks) c = 0;
                }
            }
        }
    } else if (warp == 1) {
        // =============================== MMA issuer ===============================
        if (!tc_run_fixed<STREAM, TRACE>(false, p, lx, kc, nchunks, S) && lane == 0 && !STREAM) {
            const uint32_t idesc = tc::idesc_f16(128, 2 * N);
            const uint32_t lbo_a = 128 * 16, lbo_b = 2 * N * 16;
            // descriptors differ only in the start address field (bits 0-13, 16-byte units): add offsets to a base
            const uint64_t a_base = tc::smem_desc(tc::smem_u32(ring), lbo_a, 128);
            const uint64_t b_base = tc::smem_desc(tc::smem_u32(wsm), lbo_b, 128);
            const uint32_t a_chunk = chunk_bytes >> 4, a_k16 = (2 * lbo_a) >> 4;
            const uint32_t b_chunk = ((uint32_t)(kc / 8) * lbo_b) >> 4, b_k16 = (2 * lbo_b) >> 4;
            const int crot = role.rot % nchunks;
            const int nk32 = kc / 32;
            uint32_t st = 0, ph = 0, a_ofs = 0;
            bool ok = true;
            for (int tick = 0; tick < nticks && ok; ++tick) {
                if (!tc_wait_mbar(bar_tempty, (tick & 1) ^ 1, p, 3)) break;
                tc::tc_fence_after();
                int c = crot;
                uint32_t b_ofs = (uint32_t)crot * b_chunk;
                uint32_t acc = 0;
                for (int c0 = 0; c0 < nchunks; ++c0) {
                    if (!tc_wait_mbar(bar_full + 8 * st, ph, p, 4)) { ok = false; break; }
                    if (TRACE && p.trace_mode == 1) tc_trace(p, tick, c0 & 7);
                    else if (TRACE && p.trace_mode == 0) tc_trace(p, tick, c0 == 0 ? 2 : 3);
                    tc::tc_fence_after();
                    uint64_t ad = a_base + a_ofs, bd = b_base + b_ofs;
#pragma unroll 1
                    for (int k2 = 0; k2 < ((p.debug_skip & 1) ? 0 : nk32); ++k2) {
                        tc::mma_f16(tmem_base, ad, bd, idesc, acc);
                        tc::mma_f16(tmem_base, ad + a_k16, bd + b_k16, idesc, 1u);
                        acc = 1u;
                        ad += 2 * a_k16; bd += 2 * b_k16;
                    }
                    tc::mma_commit(bar_empty + 8 * st);       // ring slot free once these MMAs have read it
                    a_ofs += a_chunk; b_ofs += b_chunk;
                    if (++st == (uint32_t)S) { st = 0; ph ^= 1; a_ofs = 0; }
                    if (++c == nchunks) { c = 0; b_ofs = 0; }
                }
                if (ok) tc::mma_commit(bar_tfull);            // accumulators of this tick complete
            }
        }
    } else {
        // =============================== epilogue ===============================
        const uint32_t trow = tmem_base + ((uint32_t)(32 * q) << 16);
        // with one round the tick-invariant addends (bias, context part of layer 1's input) live in registers; with two
        // rounds that would spill (a 320-thread CTA gets 168 registers per thread), so they are re-read every tick
        // (the same few lines each time: L1 hits)
        constexpr bool KEEP = NR == 1;
        constexpr int NRK = KEEP ? NR : 1;
        // tick-invariant addends, ONE array for both uses (registers are what this kernel is short of): the bias of an IN
        // CTA, the context part sum_k Q_k[z_k][u] of layer 1's input contraction of a REC CTA of layer 1
        float4 inv4[NRK][OWN][2];
#pragma unroll
        for (int r = 0; r < NRK; ++r)
#pragma unroll
            for (int i = 0; i < OWN; ++i) inv4[r][i][0] = inv4[r][i][1] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (KEEP && role.kind == KLM_TC_IN) {
#pragma unroll
            for (int r = 0; r < NRK; ++r)
#pragma unroll
                for (int i = 0; i < OWN; ++i)
                    if (i < rd(r).nown) {
                        const int u = 2 * (role.first_pair + rd(r).pb + rd(r).own_first + i);
                        inv4[r][i][0] = __ldg(p.bias[l] + u); inv4[r][i][1] = __ldg(p.bias[l] + u + 1);
                    }
        }
        // REC layer 1: first character index, number of real (non-padding) positions, and the context part of the
        // input contraction sum_k Q_k[z_k][u], which does not change over the ticks
        int x_next = 0, nreal_s = 0x7fffffff;
        int zk[KLM_MAX_CTX];
#pragma unroll
        for (int k = 0; k < KLM_MAX_CTX; ++k) zk[k] = 0;
        if (is_rec && l == 0 && s < p.B) {
            if (len > 0) x_next = __ldg(p.chars + o0);
            if (p.n_real) nreal_s = __ldg(p.n_real + s);
#pragma unroll
            for (int k = 0; k < KLM_MAX_CTX; ++k)
                if (k < p.Nc) {
                    int z = __ldg(p.ctx + s * p.Nc + k);
                    z = min(max(z, 0), p.ctx_rows[k] - 1);
                    zk[k] = z;
#pragma unroll
                    for (int r = 0; r < (KEEP ? NRK : 0); ++r)
#pragma unroll
                        for (int i = 0; i < OWN; ++i)
                            if (i < rd(r).nown) {
                                const int u = 2 * (role.first_pair + rd(r).pb + rd(r).own_first + i);
#pragma unroll
                                for (int e = 0; e < 2; ++e) {
                                    const float4 b = __ldg(p.Q[k] + (size_t)z * W + u + e);
                                    inv4[r][i][e].x += b.x; inv4[r][i][e].y += b.y; inv4[r][i][e].z += b.z; inv4[r][i][e].w += b.w;
                                }
                            }
                }
        }
        for (int tick = 0; tick < nticks; ++tick) {
            // the producer thread polls the grid barrier; its mbarrier arrive (release.cta) hands the
            // acquired view of the other CTAs' writes (Z, logits) on to the epilogue threads
            int ok = tc_wait_mbar(bar_go, tick & 1, p, 5) ? 1 : 0;
            ok = __all_sync(0xffffffffu, ok);
            if (!ok) break;
            const int t = tick - sigma;
            const bool active = s < p.B && t >= 0 && t < len;
            const int par_out = (tick + 1) & 1;                                         // hidden vectors: always one tick
            const int nz = p.zmask + 1, zs_in = (tick + 1 - p.zlag) & p.zmask, zs_out = (tick + 1) & p.zmask;   // Z, logits

            // REC layer 1: the character index was fetched a tick ahead
            int x = 0;
            bool real = true;
            if (is_rec && l == 0 && active) {
                x = min(max(x_next, 0), p.V - 1);
                if (t + 1 < len) x_next = __ldg(p.chars + o0 + t + 1);
                real = t < nreal_s;
            }
            // ---- addends of one round (fetched while the MMAs run / while the previous round is finished) ----
            auto load_add = [&](const int r, float4 (&add4)[OWN][2]) {
#pragma unroll
                for (int i = 0; i < OWN; ++i) add4[i][0] = add4[i][1] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (!(is_rec && active)) return;
                const int pr0 = role.first_pair + rd(r).pb + rd(r).own_first;
                if (l == 0) {
                    // fused char + context embedding gather / concat / input contraction (rating.py:103-115,125);
                    // the context part is tick-invariant (qsum)
#pragma unroll
                    for (int i = 0; i < OWN; ++i)
                        if (i < rd(r).nown) {
                            const int u = 2 * (pr0 + i);
                            // (one 256-bit load per pair and table: tc_ldg_pair)
                            float4 a0, a1;
                            tc_ldg_pair(p.P + (size_t)x * W + u, a0, a1);
                            if (real && KEEP) {
                                const float4 b0 = inv4[KEEP ? r : 0][i][0], b1 = inv4[KEEP ? r : 0][i][1];
                                a0.x += b0.x; a0.y += b0.y; a0.z += b0.z; a0.w += b0.w;
                                a1.x += b1.x; a1.y += b1.y; a1.z += b1.z; a1.w += b1.w;
                            } else {                         // padding position: context 0 (:1141-1143)
#pragma unroll
                                for (int k = 0; k < KLM_MAX_CTX; ++k)
                                    if (k < p.Nc) {
                                        float4 b0, b1;
                                        tc_ldg_pair(p.Q[k] + (real ? (size_t)zk[k] * W : (size_t)0) + u, b0, b1);
                                        a0.x += b0.x; a0.y += b0.y; a0.z += b0.z; a0.w += b0.w;
                                        a1.x += b1.x; a1.y += b1.y; a1.z += b1.z; a1.w += b1.w;
                                    }
                            }
                            add4[i][0] = a0; add4[i][1] = a1;
                        }
                } else {
                    // Z is [layer][parity][unit][64 streams] float4: the lanes of a warp (consecutive streams) touch 512
                    // contiguous bytes per unit, on the IN side (stores) as well as here
                    const float4 *zrow = reinterpret_cast<const float4 *>(p.Z) + (size_t)(l * nz + zs_in) * W * 64 + s;
#pragma unroll
                    for (int i = 0; i < OWN; ++i)
                        if (i < rd(r).nown) {
                            const int u = 2 * (pr0 + i);
                            add4[i][0] = __ldcg(zrow + (size_t)u * 64); add4[i][1] = __ldcg(zrow + (size_t)(u + 1) * 64);
                        }
                }
            };
            float4 add4[OWN][2];
            load_add(0, add4);

            // ---- accumulators ----
            ok = tc_wait_mbar(bar_tfull, tick & 1, p, 6) ? 1 : 0;
            ok = __all_sync(0xffffffffu, ok);
            if (!ok) break;
            if (TRACE && tid == 64 && (p.trace_mode == 0 || p.trace_mode == 4)) tc_trace(p, tick, p.trace_mode == 4 ? 0 : 4);
            if (early && tid == 64) {
                // the operand of this tick has been read.  What this CTA wrote in the previous tick: one bulk store that this
                // thread has waited for (pub == 2), else generic stores of all epilogue threads, ordered in front of this
                // release by the bar.sync that ended the previous tick
                if (zpub && !p.pub_release) tc::red_relaxed_add(p.barrier, 1u);
                else tc::red_release_add(p.barrier, 1u);
            }
            tc::tc_fence_after();
#pragma unroll
            for (int r = 0; r < NR; ++r) {
                const TcRound R = rd(r);
                const bool last_round = r == NR - 1 || (r + 1) * RP >= NP;      // (CTA-uniform: slices shorter than NR rounds)
                if (r > 0) tc_bar_epi();                      // the hand-over buffer is reused: round r-1 has been read
                float acc[OWN][8];
#pragma unroll
                for (int i = 0; i < OWN; ++i)
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
                if constexpr (OWN <= 2 && !STREAM) {
                // The TMEM loads of a round go out in two batches with ONE wait each (a wait per pair serialised four
                // TMEM round trips): first the pairs that the other half finishes -- this row's part of them travels
                // through shared memory (both directions at once) --, then the own pairs (hi rows: hi.hi + hi.lo columns;
                // lo rows: lo.hi columns).  Nothing may touch the destination registers between a load and its wait.
                {
                    float v[OWN][8], b[OWN][8];
#pragma unroll
                    for (int i = 0; i < OWN; ++i)
                        if (i < R.nsend) {
                            const int pr = R.pb + R.send_first + i;
                            tc::tmem_ld8(trow + 8 * pr, v[i]);
                            if (half == 0) tc::tmem_ld8(trow + N + 8 * pr, b[i]);
                        }
                    tc::tmem_ld_wait();
                    float *xs = xchg + (half ? 0 : KLM_TC1_MAX_PAIRS / 2 * 8 * 64);
#pragma unroll
                    for (int i = 0; i < OWN; ++i)
                        if (i < R.nsend) {
                            const int xi = R.send_first + i - R.oth_first;
#pragma unroll
                            for (int j = 0; j < 8; ++j) xs[(xi * 8 + j) * 64 + s] = half == 0 ? v[i][j] + b[i][j] : v[i][j];
                        }
                }
                {
                    float b[OWN][8];
#pragma unroll
                    for (int i = 0; i < OWN; ++i)
                        if (i < R.nown) {
                            const int pr = R.pb + R.own_first + i;
                            tc::tmem_ld8(trow + 8 * pr, acc[i]);
                            if (half == 0) tc::tmem_ld8(trow + N + 8 * pr, b[i]);
                        }
                    tc::tmem_ld_wait();
                    if (half == 0) {
#pragma unroll
                        for (int i = 0; i < OWN; ++i)
                            if (i < R.nown) {
#pragma unroll
                                for (int j = 0; j < 8; ++j) acc[i][j] += b[i][j];
                            }
                    }
                }
                } else {
                // own pairs: this row's part (hi rows: hi.hi + hi.lo columns; lo rows: lo.hi columns)
#pragma unroll
                for (int i = 0; i < OWN; ++i)
                    if (i < R.nown) {
                        const int pr = R.pb + R.own_first + i;
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
                    float *xs = xchg + (half ? 0 : KLM_TC1_MAX_PAIRS / 2 * 8 * 64);
#pragma unroll
                    for (int i = 0; i < OWN; ++i)
                        if (i < R.nsend) {
                            const int pr = R.pb + R.send_first + i, xi = R.send_first + i - R.oth_first;
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
                }
                if (last_round) {
                    tc::tc_fence_before();
                    tc::mbar_arrive(bar_tempty);                  // this thread is done reading TMEM
                }
                if (TRACE && tid == 64 && p.trace_mode == 4) tc_trace(p, tick, 1 + 3 * r);      // epilogue phases: round drained
                tc_bar_epi();
                {
                    const float *xr = xchg + (half ? KLM_TC1_MAX_PAIRS / 2 * 8 * 64 : 0);
#pragma unroll
                    for (int i = 0; i < OWN; ++i)
                        if (i < R.nown) {
                            const int xi = R.own_first + i - R.my_first;
#pragma unroll
                            for (int j = 0; j < 8; ++j) acc[i][j] = (acc[i][j] + xr[(xi * 8 + j) * 64 + s]) * role.scale;   // (operand scaling undone)
                        }
                    // resident form: the hand-over buffer aliases the ring: order these generic accesses before the
                    // next bulk copies
                    if (!STREAM) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                }
                if (TRACE && tid == 64 && p.trace_mode == 4) tc_trace(p, tick, 2 + 3 * r);      // ... hand-over read

                // ---- finish the round ----
                if (TRACE && last_round && tid == 64 && p.trace_mode == 0) tc_trace(p, tick, 5);
                const int pr0 = role.first_pair + R.pb + R.own_first;        // first own pair of the round (layer-wide index)
                if (is_rec) {
                    // LSTM cell for all owned units, staged so that the MUFU chains of different units overlap
                    float e1[OWN][2], e2[OWN][2], e3[OWN][2], e4[OWN][2], ag[OWN][2];
#pragma unroll
                    for (int i = 0; i < OWN; ++i)
#pragma unroll
                        for (int e = 0; e < 2; ++e) {
                            if (i >= R.nown) continue;             // (warp-uniform)
                            const float4 a = add4[i][e];
                            ag[i][e] = acc[i][4 * e + 2] + a.z;
                            e1[i][e] = tc_exp(-tc_clamp(acc[i][4 * e] + a.x));
                            e2[i][e] = tc_exp(-tc_clamp(acc[i][4 * e + 1] + a.y));
                            e3[i][e] = tc_exp(-2.f * tc_clamp(ag[i][e]));
                            e4[i][e] = tc_exp(-tc_clamp(acc[i][4 * e + 3] + a.w));
                        }
#pragma unroll
                    for (int i = 0; i < OWN; ++i)
#pragma unroll
                        for (int e = 0; e < 2; ++e) {
                            if (i >= R.nown) continue;
                            float num, den;
                            klm_tanh_parts(ag[i][e], e3[i][e], num, den);      // (small arguments: no cancellation)
                            const float ig = __fdividef(num, (1.f + e1[i][e]) * den);
                            const float f = __fdividef(1.f, 1.f + e2[i][e]);
                            e1[i][e] = fmaf(f, c_reg[r][i][e], ig);                    // new cell state
                            e2[i][e] = tc_exp(-2.f * tc_clamp(e1[i][e]));
                        }
#pragma unroll
                    for (int i = 0; i < OWN; ++i)
#pragma unroll
                        for (int e = 0; e < 2; ++e) {
                            if (i >= R.nown) continue;
                            float num, den;
                            klm_tanh_parts(e1[i][e], e2[i][e], num, den);
                            const float hn = __fdividef(num, (1.f + e4[i][e]) * den);
                            if (active) { c_reg[r][i][e] = e1[i][e]; h_reg[r][i][e] = hn; }
                        }
                    // the new hidden vector as fp16 hi/lo in the operand layout of its readers
                    if (role.pub) {
                        // slice = whole groups of 8 units: stage [group][128 rows][8]; whole 2 KB blocks are written below
#pragma unroll
                        for (int i = 0; i < OWN; ++i)
                            if (i < R.nown) {
                                const int ul = 2 * (R.pb + R.own_first + i);
                                __half hi0, lo0, hi1, lo1;
                                tc_split(h_reg[r][i][0], hi0, lo0); tc_split(h_reg[r][i][1], hi1, lo1);
                                __half *dst = pubbuf + (ul / 8) * (128 * 8) + (ul & 7);
                                *reinterpret_cast<__half2 *>(dst + s * 8) = __halves2half2(hi0, hi1);
                                *reinterpret_cast<__half2 *>(dst + (64 + s) * 8) = __halves2half2(lo0, lo1);
                            }
                    } else {
                        __half *hx_out = p.hx + ((size_t)(l * 2 + par_out) * (W / 8)) * (128 * 8);
#pragma unroll
                        for (int i = 0; i < OWN; ++i)
                            if (i < R.nown) {
                                const int u = 2 * (pr0 + i);
                                __half hi0, lo0, hi1, lo1;
                                tc_split(h_reg[r][i][0], hi0, lo0); tc_split(h_reg[r][i][1], hi1, lo1);
                                __half *dst = hx_out + (size_t)(u / 8) * (128 * 8) + (u & 7);
                                *reinterpret_cast<__half2 *>(dst + s * 8) = __halves2half2(hi0, hi1);
                                *reinterpret_cast<__half2 *>(dst + (64 + s) * 8) = __halves2half2(lo0, lo1);
                            }
                    }
                } else if (role.kind == KLM_TC_IN) {
                    // a CTA's slice of Z ([unit][64 streams] float4) is one contiguous block of 2 KB per pair: with
                    // role.pub == 2 it is staged in shared memory and leaves as one bulk async store (below), like the
                    // hidden-vector slice of a REC CTA
                    float4 *zrow = zpub ? zstage - (ptrdiff_t)(2 * role.first_pair) * 64 + s
                                                 : reinterpret_cast<float4 *>(p.Z) + (size_t)(l * nz + zs_out) * W * 64 + s;
#pragma unroll
                    for (int i = 0; i < OWN; ++i)
                        if (i < R.nown) {
                            const int u = 2 * (pr0 + i);
#pragma unroll
                            for (int e = 0; e < 2; ++e) {
                                const float4 b = KEEP ? inv4[KEEP ? r : 0][i][e] : __ldg(p.bias[l] + u + e);
                                const float4 v = make_float4(acc[i][4 * e] + b.x, acc[i][4 * e + 1] + b.y,
                                                             acc[i][4 * e + 2] + b.z, acc[i][4 * e + 3] + b.w);
                                if (zpub) zrow[(size_t)(u + e) * 64] = v;
                                else __stcg(zrow + (size_t)(u + e) * 64, v);
                            }
                        }
                } else {   // PROJ: logits of vocabulary rows 8*(first_pair + pair) ..
                    // logits are [parity][Vp/4][64 streams] float4: coalesced over the streams of a warp (read by the LSE stage)
                    float4 *lrow = zpub ? zstage - (ptrdiff_t)(2 * role.first_pair) * 64 + s
                                                 : reinterpret_cast<float4 *>(p.logits) + (size_t)zs_out * (p.Vp / 4) * 64 + s;
#pragma unroll
                    for (int i = 0; i < OWN; ++i)
                        if (i < R.nown) {
                            const int g8 = pr0 + i;
                            const float4 v0 = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
                            const float4 v1 = make_float4(acc[i][4], acc[i][5], acc[i][6], acc[i][7]);
                            if (zpub) { lrow[(size_t)(2 * g8) * 64] = v0; lrow[(size_t)(2 * g8 + 1) * 64] = v1; }
                            else { __stcg(lrow + (size_t)(2 * g8) * 64, v0); __stcg(lrow + (size_t)(2 * g8 + 1) * 64, v1); }
                        }
                }
                if (TRACE && tid == 64 && p.trace_mode == 4) tc_trace(p, tick, 3 + 3 * r);      // ... round finished
                if (last_round) break;
                // addends of the next round, fetched BEHIND this round's cell and staging (in front of them the issue of the
                // scattered table-row loads of layer 1 alone was ~1 800 clk on the recurrence); they are in flight while the
                // next round drains its accumulators and exchanges the hi / lo parts
                if (r + 1 < NR) load_add(r + 1, add4);
            }

            // ---- hand over ----
            if (is_rec) {
                if (role.pub) {
                    __half *gdst = p.hx + ((size_t)(l * 2 + par_out) * (W / 8) + role.first_pair / 4) * (128 * 8);
                    const int n16 = (NP / 4) * 128;                       // 16-byte pieces of the slice
                    if (role.pub == 2) {
                        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // staging writes -> bulk store reads
                        if (TRACE && tid == 64 && p.trace_mode == 0) tc_trace(p, tick, 6);
                        tc_bar_epi();
                        if (tid == 64) {
                            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                                         ::"l"(gdst), "r"(tc::smem_u32(pubbuf)), "r"((uint32_t)n16 * 16u) : "memory");
                            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                            asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // writes complete (not just read)
                            // everything this CTA hands over in a tick went through that one bulk store, and it has been
                            // performed: a release (a gpu-scope fence in front of the add, ~900 clk on the recurrence:
                            // scripts/ubench/pingpong.cu) has nothing left to order
                            if (p.fault && role.rot == 0 && tick == 2) { }          // (test hook: the watchdog must end the launch)
                            else if (p.pub_release) tc::red_release_add(p.barrier, 1u);
                            else tc::red_relaxed_add(p.barrier, 1u);
                            if (TRACE && (p.trace_mode == 0 || p.trace_mode == 4)) tc_trace(p, tick, 7);
                        }
                        continue;
                    }
                    tc_bar_epi();
                    const uint4 *sp = reinterpret_cast<const uint4 *>(pubbuf);
                    uint4 *gp = reinterpret_cast<uint4 *>(gdst);
                    for (int i = tid - 64; i < n16; i += 256) __stcg(gp + i, sp[i]);
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");       // staging reads -> next bulk copies into the ring
                }
                tc::fence_proxy_async();             // generic global writes -> bulk (async proxy) reads of other CTAs
            }
            if (zpub) {
                // IN / PROJ: the staged slice of Z / of the logits as one bulk async store; once it has been performed a
                // relaxed add signals it (no generic global store of this CTA is outstanding)
                float4 *gdst = role.kind == KLM_TC_IN
                    ? reinterpret_cast<float4 *>(p.Z) + ((size_t)(l * nz + zs_out) * W + 2 * role.first_pair) * 64
                    : reinterpret_cast<float4 *>(p.logits) + ((size_t)zs_out * (p.Vp / 4) + 2 * role.first_pair) * 64;
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // staging writes -> bulk store reads
                if (TRACE && tid == 64 && p.trace_mode == 0) tc_trace(p, tick, 6);
                tc_bar_epi();
                if (tid == 64) {
                    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                                 ::"l"(gdst), "r"(tc::smem_u32(zstage)), "r"((uint32_t)NP * 2048u) : "memory");
                    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
                    if (!early) {
                        if (p.pub_release) tc::red_release_add(p.barrier, 1u);
                        else tc::red_relaxed_add(p.barrier, 1u);
                    }
                    if (guard) tc::mbar_arrive(bar_epid);
                    if (TRACE && (p.trace_mode == 0 || p.trace_mode == 4)) tc_trace(p, tick, 7);
                }
                continue;
            }
            if (TRACE && tid == 64 && p.trace_mode == 0) tc_trace(p, tick, 6);
            tc_bar_epi();
            if (tid == 64) {
                if (!early) tc::red_release_add(p.barrier, 1u);
                if (guard) tc::mbar_arrive(bar_epid);
                if (TRACE && (p.trace_mode == 0 || p.trace_mode == 4)) tc_trace(p, tick, 7);
            }
        }
        // ---- write the final state back ----
        if (is_rec && s < p.B) {
            float *st = p.state + ((size_t)p.slots_out[s] * p.D + l) * 2 * W;
#pragma unroll
            for (int r = 0; r < NR; ++r)
#pragma unroll
                for (int i = 0; i < OWN; ++i)
                    if (i < rd(r).nown) {
                        const int u = 2 * (role.first_pair + rd(r).pb + rd(r).own_first + i);
                        st[u] = h_reg[r][i][0]; st[u + 1] = h_reg[r][i][1];
                        st[W + u] = c_reg[r][i][0]; st[W + u + 1] = c_reg[r][i][1];
                    }
        }
    }

    tc::tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        __syncwarp();
        tc::tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

template <bool STREAM, bool TRACE>
static __device__ __forceinline__ void tc_seq_body(const KlmTcParams &p, unsigned char *smem) {
    const KlmTcRole role = p.roles[blockIdx.x];
    if (role.kind == KLM_TC_LSE) { tc_lse_role(p, role, smem); return; }
    if (STREAM) {
        // streamed-weights form: slices of up to 24 pairs in up to three rounds of 8 (2 pairs per thread and round)
        if (role.kind == KLM_TC_REC) tc_seq_role<3, STREAM, TRACE, 2, true>(p, role, smem);
        else tc_seq_role<3, STREAM, TRACE, 2, false>(p, role, smem);
    } else if (role.kind == KLM_TC_REC) {
        if (role.npairs <= 8) tc_seq_role<1, STREAM, TRACE, 2, true>(p, role, smem);
        else tc_seq_role<1, STREAM, TRACE, KLM_TC_OWN, true>(p, role, smem);
    } else {
        tc_seq_role<1, STREAM, TRACE, KLM_TC_OWN, false>(p, role, smem);
    }
}

__global__ void __launch_bounds__(KLM_TC_THREADS, 1) klm_seq_tc(const __grid_constant__ KlmTcParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    tc_seq_body<false, false>(p, smem);
}
__global__ void __launch_bounds__(KLM_TC_THREADS, 1) klm_seq_tc_traced(const __grid_constant__ KlmTcParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    tc_seq_body<false, true>(p, smem);
}

// Streamed-weights form: the B images stay in global memory (L2) and travel through the ring with the A chunks.
__global__ void __launch_bounds__(KLM_TC_THREADS, 1) klm_seq_tcs(const __grid_constant__ KlmTcParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    tc_seq_body<true, false>(p, smem);
}
__global__ void __launch_bounds__(KLM_TC_THREADS, 1) klm_seq_tcs_traced(const __grid_constant__ KlmTcParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    tc_seq_body<true, true>(p, smem);
}

cudaError_t klm_launch_seq_tc(const KlmTcParams &p, int grid, size_t smem, cudaStream_t stream) {
    void *args[] = { const_cast<KlmTcParams *>(&p) };
    // (the attribute is per device: set it once per device and size, not at every launch of a beam step)
    static size_t smem_set[4][64] = {{0}};
    const void *fns[4] = {(const void *)klm_seq_tc, (const void *)klm_seq_tcs, (const void *)klm_seq_tc_traced,
                          (const void *)klm_seq_tcs_traced};
    const int which = (p.stream_b ? 1 : 0) + (p.trace ? 2 : 0);     // the traced instantiations only while klm_tc_trace is armed
    const void *fn = fns[which];
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64 || smem_set[which][dev] != smem) {
        cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        if (dev >= 0 && dev < 64) smem_set[which][dev] = smem;
    }
    cudaError_t e = cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(KLM_TC_THREADS), args, smem, stream);
    if (e != cudaSuccess) {
        cudaFuncAttributes fa;
        int occ = -1;
        if (cudaFuncGetAttributes(&fa, fn) == cudaSuccess) {
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, KLM_TC_THREADS, smem);
            fprintf(stderr, "klm_launch_seq_tc: %s; grid %d, %d threads, %zu B dynamic + %zu B static shared memory (limit %d), "
                    "%d registers, %zu B local, max threads %d, blocks per SM %d\n", cudaGetErrorString(e), grid, KLM_TC_THREADS,
                    smem, fa.sharedSizeBytes, fa.maxDynamicSharedSizeBytes, fa.numRegs, fa.localSizeBytes, fa.maxThreadsPerBlock, occ);
        }
    }
    return e;
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
