// Dataflow form of the persistent stacked-LSTM sequence kernel on the 5th-generation tensor cores
// (tcgen05 + TMEM), for 9..64 concurrent streams (sequence mode) and for beam steps of up to 256 hypotheses
// (step mode).  Replaces the per-window Session.run of Rater.rate (rating.py:516) and the per-step
// predict_on_batch of Rater.predict / rate_best (:631, :789-852).
//
// Same GEMM decomposition as klm_seq_tc.cu (every CTA owns N <= 96 columns of ONE weight matrix, resident in
// shared memory as an fp16 hi/lo B image; per tick D[128 x 2N] = A[128 x W] . B[2N x W]^T with A = one layer's
// hidden vector for 64 streams, rows 0..63 hi part, 64..127 lo part; pre-activation = hi.hi + hi.lo + lo.hi), but
// NO grid-wide barrier and no fixed wavefront: every hand-over has its own flag or counter, every CTA runs its
// ticks as soon as its inputs are there, and all stages work on position `tick`:
//
//   REC l   h_l[t] = cell(R_l . h_l[t-1] + (l = 0 ? P[x_t] + ctx : Z_l[t]))     needs: all slices of h_l[t-1] (flags),
//                                                                                Z_l[t] complete (counter)
//   IN  l   Z_l[t] = K_l . h_{l-1}[t] + b_l                                      needs: all slices of h_{l-1}[t] (flags)
//   PROJ    logits[t] = E . h_{D-1}[t]                                            needs: all slices of h_{D-1}[t] (flags)
//   LSE     log-softmax / target pick, per warp                                  needs: logits[t] complete (counter)
//
// * A REC CTA publishes its slice of the new hidden vector as ONE bulk async store into version (t+2) of a ring of
//   KLM_DF_R versions and then sets its flag to the version number; the producer warp of a reader polls the flags of
//   the source layer (one coalesced load per poll) and issues the bulk copy of an operand chunk as soon as the slices
//   that make up THAT chunk are there -- the fetch overlaps the stragglers, and nobody waits for CTAs it does not
//   read from (round 1: one grid barrier per tick, ~2 000 clk of propagation + every CTA waited for all 148).
// * IN / PROJ CTAs do not sit on the recurrence at all: their hi-row and lo-row partial sums go out as two arrays
//   (Z / logits parts, added by the consumer), so they need no hi<->lo hand-over, no shared-memory staging and no
//   buffer aliased on the operand ring; back-pressure (a slot of Z / of the logits is reused, a version of the
//   hidden-vector ring is overwritten) runs through counters that are polled long before they matter.
// * In a REC CTA the hand-over buffer and the publication staging alias the ring slot of the chunk that contains the
//   CTA's OWN slice: that slot is refilled only after the CTA's own flag is up, i.e. after its publication is through.
// * step mode (klm_step, beam search): one position for up to 64 * KLM_DF_MAXG hypotheses in ONE launch; tick g
//   advances hypotheses 64g..64g+63.  The initial hidden vectors of all groups are published in the prologue
//   (hx0), so every REC CTA runs its R_l . h_l products back to back without waiting for anybody, and the chain of a
//   group is REC 0 -> IN 1 -> cell of REC 1 -> IN 2 -> ... -> PROJ -> LSE (round 1: 2D+1 full ticks per 64 hypotheses).
//   Cell and hidden states travel through the state pool (slots_in / slots_out), not through registers.
// * Flags and counters are never reset: a launch works above KlmTcParams::epoch / tick_base (no memset per launch).
//
// Warp roles in a GEMM CTA (320 threads): warp 0 = producer (flag polls + bulk copies, one elected lane issues),
// warp 1 = MMA issuer, warps 2..9 = epilogue.  Operands are scaled by powers of two before the fp16 split
// (KLM_TC_HSCALE), KlmTcRole::scale undoes it on the fp32 accumulator.
#include <cstdio>
#include "klm_tc_dev.cuh"

namespace {

__device__ __forceinline__ unsigned df_ld_acquire(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void df_st_relaxed(unsigned *p, unsigned v) {
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void df_st_release(unsigned *p, unsigned v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ bool df_reached(unsigned v, unsigned target) { return (int)(v - target) >= 0; }
// Signals a slice that went out as a bulk async store the calling thread has waited for (cp.async.bulk.wait_group 0).
// mode 1 (default): gpu-scope release.  A relaxed store (mode 0) is NOT enough: the flag can be seen before the data
// (scripts/exp/df_race.py); modes 2 / 3 are experiments (proxy fence + relaxed store; delay + relaxed store).
__device__ __forceinline__ void df_signal(unsigned *flag, unsigned value, int mode) {
    if (mode == 1) df_st_release(flag, value);
    else if (mode == 4) {           // (experiment: the release as an atomic performed at the home L2 partition)
        unsigned old;
        asm volatile("atom.release.gpu.global.exch.b32 %0, [%1], %2;" : "=r"(old) : "l"(flag), "r"(value) : "memory");
    } else {
        if (mode == 2) tc::fence_proxy_async();
        else if (mode == 3) __nanosleep(300);
        df_st_relaxed(flag, value);
    }
}

// one thread waits until *w has reached `target` (wrap-safe); false = watchdog
__device__ __forceinline__ bool df_wait_word(const KlmTcParams &p, const unsigned *w, unsigned target, int code) {
    unsigned n = 0;
    while (!df_reached(df_ld_acquire(w), target))
        if (tc_spin_fail(n, p, code)) return false;
    return true;
}

// all lanes of the calling warp: wait until the words w[0 .. n) (n <= 64) have ALL reached `target`; false = watchdog
__device__ __forceinline__ bool df_wait_all(const KlmTcParams &p, const unsigned *w, int n, unsigned target, int code) {
    const int lane = threadIdx.x & 31;
    unsigned cnt = 0;
    for (;;) {
        bool ok = lane >= n || df_reached(df_ld_acquire(w + lane), target);
        if (n > 32) ok = (lane + 32 >= n || df_reached(df_ld_acquire(w + 32 + lane), target)) && ok;
        if (__all_sync(0xffffffffu, ok ? 1 : 0)) return true;
        const int bad = tc_spin_fail(cnt, p, code) ? 1 : 0;
        if (__any_sync(0xffffffffu, bad)) return false;
    }
}
// one look, no waiting
__device__ __forceinline__ bool df_check_all(const unsigned *w, int n, unsigned target) {
    const int lane = threadIdx.x & 31;
    bool ok = lane >= n || df_reached(df_ld_acquire(w + lane), target);
    if (n > 32) ok = (lane + 32 >= n || df_reached(df_ld_acquire(w + 32 + lane), target)) && ok;
    return __all_sync(0xffffffffu, ok ? 1 : 0) != 0;
}

struct DfCtx {
    uint32_t ring_a, bar_full, bar_empty, bar_tfull, bar_tempty, bar_epid, tmem_base, wsm_a, masks_a;
    int T, N, NP, crot, W, guard, nch;
    // what changes from step to step of a resident instance (else: copies of the kernel parameters)
    unsigned epoch, tick_base;
    int n_hyp, groups;
};

// Split of a CTA's NP pairs between the hi-row and the lo-row threads, and between the two warps that share a TMEM
// lane quarter (REC: the pair is finished by its owner, the other half's part arrives through shared memory).
struct DfSplit { int my_first, oth_first, own_first, nown, send_first, nsend; };
__device__ __forceinline__ DfSplit df_split(int NP, int half, int ch) {
    DfSplit o;
    const int Ph = (NP + 1) / 2;
    const int my_n = half ? NP - Ph : Ph, oth_n = half ? Ph : NP - Ph;
    o.my_first = half ? Ph : 0;
    o.oth_first = half ? 0 : Ph;
    o.own_first = o.my_first + (ch ? (my_n + 1) / 2 : 0);
    o.nown = ch ? my_n - (my_n + 1) / 2 : (my_n + 1) / 2;
    o.send_first = o.oth_first + (ch ? (oth_n + 1) / 2 : 0);
    o.nsend = ch ? oth_n - (oth_n + 1) / 2 : (oth_n + 1) / 2;
    return o;
}

// ---------------------------------------------------------------------------------------------
// Producer warp: per tick and operand chunk -- ring slot free, the slices of the chunk published, bulk copy.
template <int KC, int S, bool TRACE>
__device__ __noinline__ void df_produce(const KlmTcParams &p, const KlmTcRole &role, const DfCtx &x) {
    constexpr uint32_t CB = KC * 256;
    const uint32_t ring_a = x.ring_a, bar_full = x.bar_full, bar_empty = x.bar_empty, bar_epid = x.bar_epid, masks_a = x.masks_a;
    const int T = x.T, crot = x.crot, NCH = x.nch, guard = x.guard, nsrc = role.n_src;
    const bool is_rec = role.kind == KLM_TC_REC, step_rec = is_rec && x.groups > 0;
    const int lane = threadIdx.x & 31;
    const size_t hx_bytes = (size_t)x.W * 128 * sizeof(__half);
    const unsigned char *hx = reinterpret_cast<const unsigned char *>(p.hx) + (size_t)role.src_layer * KLM_DF_R * hx_bytes;
    const unsigned char *hx0 = reinterpret_cast<const unsigned char *>(p.hx0) + (size_t)role.src_layer * KLM_DF_MAXG * hx_bytes;
    const unsigned *flags = p.sync + (step_rec ? KLM_DF_HFLAG0 : KLM_DF_HFLAG) + role.src_layer * 64;
    const unsigned epoch = x.epoch;
    const int cfence = p.consumer_fence;
    uint32_t phbits = 0;
    bool alive = true;
    for (int tick = 0; tick < T && alive; ++tick) {
        const unsigned version = (unsigned)tick + (is_rec ? 1u : 2u);
        const unsigned target = step_rec ? epoch + 1u : epoch + version;
        const unsigned char *src = step_rec ? hx0 + (size_t)tick * hx_bytes : hx + (size_t)(version & (KLM_DF_R - 1)) * hx_bytes;
        if (guard) {     // aliased buffers not covered by the own-flag rule: wait for the epilogue (arrival 0 = its prologue)
            alive = __all_sync(0xffffffffu, tc_wait_mbar(bar_epid, (uint32_t)tick & 1u, p, 7) ? 1 : 0);
            if (!alive) break;
        }
        unsigned long long have = 0;
        int c = crot;
        for (int c0 = 0; c0 < NCH && alive; c0 += S) {
#pragma unroll
            for (int j = 0; j < S; ++j) {
                if (c0 + j < NCH && alive) {
                    alive = __all_sync(0xffffffffu, tc_wait_mbar(bar_empty + 8 * j, ((phbits >> j) & 1u) ^ 1u, p, 2) ? 1 : 0);
                    if (alive) {
                        if (tc::elect_one()) tc::mbar_arrive_expect_tx(bar_full + 8 * j, CB);
                        unsigned long long need;
                        asm volatile("ld.shared.u64 %0, [%1];" : "=l"(need) : "r"(masks_a + 8u * (uint32_t)c));
                        unsigned n = 0;
                        while ((have & need) != need) {
                            const unsigned v0 = lane < nsrc ? df_ld_acquire(flags + lane) : target;
                            unsigned long long m = __ballot_sync(0xffffffffu, df_reached(v0, target));
                            if (nsrc > 32) {
                                const unsigned v1 = lane + 32 < nsrc ? df_ld_acquire(flags + 32 + lane) : target;
                                m |= (unsigned long long)__ballot_sync(0xffffffffu, df_reached(v1, target)) << 32;
                            }
                            have |= m;
                            if ((have & need) == need) break;
                            if (p.exp & 1) __nanosleep(100);
                            int bad = tc_spin_fail(n, p, 1) ? 1 : 0;
                            if (__any_sync(0xffffffffu, bad)) { alive = false; break; }
                        }
                        if (alive) {
                            __syncwarp();
                            if (tc::elect_one()) {
                                if (TRACE && c0 + j == 0) tc_trace(p, tick, 0);
                                if (cfence) tc::fence_proxy_async();
                                tc::bulk_g2s(ring_a + j * CB, src + (size_t)c * CB, CB, bar_full + 8 * j);
                                if (TRACE && c0 + j == 0) tc_trace(p, tick, 1);
                            }
                            __syncwarp();
                            phbits ^= 1u << j;
                            c = (c + 1 == NCH) ? 0 : c + 1;
                        }
                    }
                }
            }
        }
    }
}

// MMA warp: per tick, the chunks in the order the producer fetches them (fixed per CTA: results are deterministic).
template <int KC, int S, bool TRACE>
__device__ __noinline__ void df_mma(const KlmTcParams &p, const DfCtx &x) {
    constexpr uint32_t CB = KC * 256;
    const uint32_t idesc = tc::idesc_f16(128, 2 * x.N);
    const uint32_t lbo_a = 128 * 16, lbo_b = 2 * x.N * 16;
    constexpr uint32_t slot16 = CB >> 4;
    const uint64_t a_base = tc::smem_desc(x.ring_a, lbo_a, 128);
    const uint64_t b_base = tc::smem_desc(x.wsm_a, lbo_b, 128);
    constexpr uint32_t a_k16 = (2 * 128 * 16) >> 4;
    const uint32_t b_k16 = (2 * lbo_b) >> 4, b_chunk = ((uint32_t)(KC / 8) * lbo_b) >> 4;
    const uint32_t bar_full = x.bar_full, bar_empty = x.bar_empty, bar_tfull = x.bar_tfull, bar_tempty = x.bar_tempty;
    const uint32_t tmem_base = x.tmem_base;
    const int T = x.T, crot = x.crot, NCH = x.nch;
    uint32_t phbits = 0;
    bool alive = true;
    for (int tick = 0; tick < T && alive; ++tick) {
        alive = __all_sync(0xffffffffu, tc_wait_mbar(bar_tempty, (tick & 1) ^ 1, p, 3) ? 1 : 0);
        if (!alive) break;
        tc::tc_fence_after();
        int c = crot;
        for (int c0 = 0; c0 < NCH && alive; c0 += S) {
#pragma unroll
            for (int j = 0; j < S; ++j) {
                if (c0 + j < NCH && alive) {
                    alive = __all_sync(0xffffffffu, tc_wait_mbar(bar_full + 8 * j, (phbits >> j) & 1u, p, 4) ? 1 : 0);
                    if (alive) {
                        const uint64_t ad = a_base + (uint32_t)(j * slot16);
                        const uint64_t bd = b_base + (uint32_t)c * b_chunk;
                        const bool last = c0 + j == NCH - 1;
                        if (tc::elect_one()) {
                            if (TRACE && (c0 + j == 0 || last)) tc_trace(p, tick, c0 + j == 0 ? 2 : 3);
#pragma unroll
                            for (int kk = 0; kk < KC / 16; ++kk)
                                tc::mma_f16(tmem_base, ad + kk * a_k16, bd + kk * b_k16, idesc, (c0 + j + kk) != 0 ? 1u : 0u);
                            tc::mma_commit(bar_empty + 8 * j);       // ring slot free once these MMAs have read it
                            if (last) tc::mma_commit(bar_tfull);     // accumulators of this tick complete
                        }
                        __syncwarp();
                        phbits ^= 1u << j;
                        c = (c + 1 == NCH) ? 0 : c + 1;
                    }
                }
            }
        }
    }
}

template <bool TRACE>
__device__ __forceinline__ bool df_run_loops(bool producer, const KlmTcParams &p, const KlmTcRole &role, const DfCtx &x) {
#define DF_TRY(KC_, S_)                                                     \
    if (role.kc == KC_ && role.stages == S_) {                              \
        if (producer) df_produce<KC_, S_, TRACE>(p, role, x);               \
        else df_mma<KC_, S_, TRACE>(p, x);                                  \
        return true;                                                        \
    }
    DF_TRY(128, 3) DF_TRY(64, 3) DF_TRY(64, 4) DF_TRY(128, 2) DF_TRY(128, 1) DF_TRY(64, 2) DF_TRY(64, 1)
#undef DF_TRY
    return false;
}

// ---------------------------------------------------------------------------------------------
// Epilogue of a REC CTA: accumulators -> (hi/lo hand-over) -> + addend -> LSTM cell -> new hidden slice published.
// OWN = pairs of units finished by one thread (ceil(ceil(NP / 2) / 2)).
template <int OWN, bool TRACE>
__device__ __noinline__ void df_epi_rec(const KlmTcParams &p, const KlmTcRole &role, const DfCtx &x, unsigned char *alias) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int W = p.W, NP = x.NP, N = x.N, T = x.T, l = role.layer;
    const int q = warp & 3, row = 32 * q + lane, half = row >> 6, s = row & 63, ch = (warp - 2) >> 2;
    const DfSplit R = df_split(NP, half, ch);
    const int xh = (NP + 1) / 2;                                   // pairs per direction of the hand-over buffer
    float *xchg = reinterpret_cast<float *>(alias);
    __half *pubbuf = reinterpret_cast<__half *>(alias + 2 * xh * 2048);
    const uint32_t trow = x.tmem_base + ((uint32_t)(32 * q) << 16);
    const bool step = x.groups > 0;
    const float scale = role.scale;
    const int pr0 = role.first_pair + R.own_first;                 // first own pair (layer-wide index)
    const size_t hx_slice = (size_t)W * 128;                       // halves of one version of one layer
    __half *hx_l = p.hx + (size_t)l * KLM_DF_R * hx_slice;
    unsigned *my_flag = p.sync + KLM_DF_HFLAG + l * 64 + role.idx;
    const unsigned epoch = x.epoch, tick_base = x.tick_base;
    const uint32_t pub_bytes = (uint32_t)(NP / 4) * 2048u;

    float c_reg[OWN][2], h_reg[OWN][2];
    long long o0 = 0;
    int len = 0;
#pragma unroll
    for (int i = 0; i < OWN; ++i) c_reg[i][0] = c_reg[i][1] = h_reg[i][0] = h_reg[i][1] = 0.f;

    // stages the hidden values in h_reg as fp16 hi/lo in the readers' operand layout ([group of 8 units][128 rows][8])
    auto stage = [&]() {
#pragma unroll
        for (int i = 0; i < OWN; ++i)
            if (i < R.nown) {
                const int ul = 2 * (R.own_first + i);
                __half hi0, lo0, hi1, lo1;
                tc_split(h_reg[i][0], hi0, lo0); tc_split(h_reg[i][1], hi1, lo1);
                __half *dst = pubbuf + (ul / 8) * (128 * 8) + (ul & 7);
                *reinterpret_cast<__half2 *>(dst + s * 8) = __halves2half2(hi0, hi1);
                *reinterpret_cast<__half2 *>(dst + (64 + s) * 8) = __halves2half2(lo0, lo1);
            }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");       // staging writes -> bulk store reads
    };
    // one thread: the staged slice as ONE bulk async store; returns when the writes have been performed
    auto bulk_out = [&](__half *gdst) {
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                     ::"l"(gdst), "r"(tc::smem_u32(pubbuf)), "r"(pub_bytes) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    };
    auto load_state = [&](int slot, bool with_h) {
#pragma unroll
        for (int i = 0; i < OWN; ++i) {
            c_reg[i][0] = c_reg[i][1] = 0.f;
            if (with_h) h_reg[i][0] = h_reg[i][1] = 0.f;
            if (i < R.nown && slot >= 0) {
                const int u = 2 * (pr0 + i);
                const float *st = p.state + ((size_t)slot * p.D + l) * 2 * W;
                if (with_h) { const float2 h2 = *reinterpret_cast<const float2 *>(st + u); h_reg[i][0] = h2.x; h_reg[i][1] = h2.y; }
                const float2 c2 = *reinterpret_cast<const float2 *>(st + W + u);
                c_reg[i][0] = c2.x; c_reg[i][1] = c2.y;
            }
        }
    };

    // ---- prologue: initial hidden vectors ----
    if (!step) {
        if (s < p.B) { o0 = p.offsets[s]; len = (int)(p.offsets[s + 1] - o0); }
        load_state(s < p.B ? p.slots_in[s] : -1, true);
        stage();
        tc_bar_epi();
        if (tid == 64) {
            bulk_out(hx_l + (size_t)1 * hx_slice + (size_t)(role.first_pair / 4) * (128 * 8));     // version 1
            df_signal(my_flag, epoch + 1u, p.pub_release);
        }
    } else {
        // the initial hidden values of ALL groups: loads first (they are independent), staged side by side behind the
        // hand-over buffer, one bulk store per group, ONE wait and one release for all of them
        __half *hx0_l = p.hx0 + (size_t)l * KLM_DF_MAXG * hx_slice;
        float h0[KLM_DF_MAXG][OWN][2];
#pragma unroll
        for (int g = 0; g < KLM_DF_MAXG; ++g) {
            const int hyp = 64 * g + s;
            const int slot = (g < x.groups && hyp < x.n_hyp) ? (p.step_inline ? p.step.slots_in[hyp] : p.slots_in[hyp]) : -1;
#pragma unroll
            for (int i = 0; i < OWN; ++i) {
                h0[g][i][0] = h0[g][i][1] = 0.f;
                if (i < R.nown && slot >= 0) {
                    const float2 h2 = *reinterpret_cast<const float2 *>(p.state + ((size_t)slot * p.D + l) * 2 * W + 2 * (pr0 + i));
                    h0[g][i][0] = h2.x; h0[g][i][1] = h2.y;
                }
            }
        }
#pragma unroll
        for (int g = 0; g < KLM_DF_MAXG; ++g)
            if (g < x.groups) {
                __half *pb = pubbuf + (size_t)g * (pub_bytes / 2);
#pragma unroll
                for (int i = 0; i < OWN; ++i)
                    if (i < R.nown) {
                        const int ul = 2 * (R.own_first + i);
                        __half hi0, lo0, hi1, lo1;
                        tc_split(h0[g][i][0], hi0, lo0); tc_split(h0[g][i][1], hi1, lo1);
                        __half *dst = pb + (ul / 8) * (128 * 8) + (ul & 7);
                        *reinterpret_cast<__half2 *>(dst + s * 8) = __halves2half2(hi0, hi1);
                        *reinterpret_cast<__half2 *>(dst + (64 + s) * 8) = __halves2half2(lo0, lo1);
                    }
            }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        tc_bar_epi();
        if (tid == 64) {
            for (int g = 0; g < x.groups; ++g)
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                             ::"l"(hx0_l + (size_t)g * hx_slice + (size_t)(role.first_pair / 4) * (128 * 8)),
                               "r"(tc::smem_u32(pubbuf) + (uint32_t)g * pub_bytes), "r"(pub_bytes) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
            unsigned *f0 = p.sync + KLM_DF_HFLAG0 + l * 64 + role.idx;
            df_signal(f0, epoch + 1u, p.pub_release);
        }
    }
    tc_bar_epi();                                                    // (staging free again before the ring is used)
    if (tid == 64 && x.guard) tc::mbar_arrive(x.bar_epid);
    if (TRACE && tid == 64) tc_trace(p, p.trace_t0 + KLM_TC_TRACE_TICKS - 1, 2);

    int x_next = 0, nreal_s = 0x7fffffff;
    if (l == 0 && !step && s < p.B) {
        if (len > 0) x_next = __ldg(p.chars + o0);
        if (p.n_real) nreal_s = __ldg(p.n_real + s);
    }
    bool alive = true;
    for (int tick = 0; tick < T; ++tick) {
        const int hyp = step ? 64 * tick + s : s;
        const bool active = step ? hyp < x.n_hyp : (s < p.B && tick < len);
        const long long pos = step ? hyp : o0 + tick;
        int slot_out = 0;
        if (step) {
            load_state(active ? (p.step_inline ? p.step.slots_in[hyp] : p.slots_in[hyp]) : -1, false);
            if (active) slot_out = p.step_inline ? p.step.slots_out[hyp] : p.slots_out[hyp];   // (fetched now, needed behind the cell)
        }
        // back-pressure on the version ring: version tick+2 overwrites version tick+2-R, read by the IN / PROJ CTAs in tick
        // tick-R.  Polled here, needed only at the very end of the tick.
        bool bp_ok = true;
        if (warp == 2 && tick >= KLM_DF_R && role.n_readers > 0)
            bp_ok = df_check_all(p.sync + KLM_DF_HRD + l * 64, role.n_readers, tick_base + (unsigned)(tick - KLM_DF_R + 1));

        // ---- addends: layer 0 = fused char + context embedding gather (tables P, Q: rating.py:103-115,125); else Z ----
        float4 add4[OWN][2];
#pragma unroll
        for (int i = 0; i < OWN; ++i) add4[i][0] = add4[i][1] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (l == 0) {
            if (active) {
                int xc;
                bool real = true;
                if (step) xc = p.step_inline ? p.step.chars[hyp] : __ldg(p.chars + hyp);
                else {
                    xc = x_next;
                    if (tick + 1 < len) x_next = __ldg(p.chars + o0 + tick + 1);
                    real = tick < nreal_s;
                }
                xc = min(max(xc, 0), p.V - 1);
#pragma unroll
                for (int i = 0; i < OWN; ++i)
                    if (i < R.nown) {
                        const int u = 2 * (pr0 + i);
#pragma unroll
                        for (int e = 0; e < 2; ++e) {
                            float4 a = __ldg(p.P + (size_t)xc * W + u + e);
#pragma unroll
                            for (int k = 0; k < KLM_MAX_CTX; ++k)
                                if (k < p.Nc) {
                                    int z = !real ? 0 : (step && p.step_inline) ? p.step.ctx[hyp * p.Nc + k]
                                                                                : __ldg(p.ctx + (size_t)hyp * p.Nc + k);   // padding: context 0 (:1141-1143)
                                    z = min(max(z, 0), p.ctx_rows[k] - 1);
                                    const float4 b = __ldg(p.Q[k] + (size_t)z * W + u + e);
                                    a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
                                }
                            add4[i][e] = a;
                        }
                    }
            }
        } else {
            // Z_l[tick] complete?  Every IN CTA of the layer has its own progress word (in steady state the IN stage is ahead)
            if (alive) alive = df_wait_all(p, p.sync + KLM_DF_ZPROG + l * 64, role.n_zprod, tick_base + (unsigned)tick + 1u, 8);
            const float4 *za = reinterpret_cast<const float4 *>(p.Z) + ((size_t)(l * KLM_DF_NZ + (tick & (KLM_DF_NZ - 1))) * 2) * W * 64 + s;
            const float4 *zb = za + (size_t)W * 64;
#pragma unroll
            for (int i = 0; i < OWN; ++i)
                if (i < R.nown) {
                    const int u = 2 * (pr0 + i);
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const float4 a = __ldcg(za + (size_t)(u + e) * 64), b = __ldcg(zb + (size_t)(u + e) * 64);
                        add4[i][e] = make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
                    }
                }
        }

        // ---- accumulators ----
        {
            int ok = alive ? (tc_wait_mbar(x.bar_tfull, tick & 1, p, 6) ? 1 : 0) : 1;
            alive = alive && __all_sync(0xffffffffu, ok);
        }
        if (TRACE && tid == 64) tc_trace(p, tick, 4);
        tc::tc_fence_after();
        float acc[OWN][8];
#pragma unroll
        for (int i = 0; i < OWN; ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
        if constexpr (OWN <= 0) {        // (batched form as in klm_seq_tc: no gain on a beam step, and it costs the registers the step server needs)
        // two batches of TMEM loads with one wait each: the pairs the other half finishes (this row's part goes over through
        // shared memory, both directions at once), then the own pairs (hi rows: hi.hi + hi.lo columns; lo rows: lo.hi columns)
        {
            float v[OWN][8], b[OWN][8];
#pragma unroll
            for (int i = 0; i < OWN; ++i)
                if (i < R.nsend) {
                    const int pr = R.send_first + i;
                    tc::tmem_ld8(trow + 8 * pr, v[i]);
                    if (half == 0) tc::tmem_ld8(trow + N + 8 * pr, b[i]);
                }
            tc::tmem_ld_wait();
            float *xs = xchg + (half ? 0 : xh * 8 * 64);
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
                    const int pr = R.own_first + i;
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
                const int pr = R.own_first + i;
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
        // the other half's pairs: this row's part goes over through shared memory (both directions at once)
        {
            float *xs = xchg + (half ? 0 : xh * 8 * 64);
#pragma unroll
            for (int i = 0; i < OWN; ++i)
                if (i < R.nsend) {
                    const int pr = R.send_first + i, xi = R.send_first + i - R.oth_first;
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
        tc::tc_fence_before();
        tc::mbar_arrive(x.bar_tempty);                               // this thread is done reading TMEM
        tc_bar_epi();
        {
            const float *xr = xchg + (half ? xh * 8 * 64 : 0);
#pragma unroll
            for (int i = 0; i < OWN; ++i)
                if (i < R.nown) {
                    const int xi = R.own_first + i - R.my_first;
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[i][j] += xr[(xi * 8 + j) * 64 + s];
                }
        }
        if (TRACE && tid == 64) tc_trace(p, tick, 5);

        // ---- LSTM cell (Keras LSTMCell, gate order i|f|c|o, exact logistic: rating.py:133-134), 8 MUFU operations per unit,
        //      staged so that the chains of different units overlap ----
        {
            float e1[OWN][2], e2[OWN][2], e3[OWN][2], e4[OWN][2], ag[OWN][2];
#pragma unroll
            for (int i = 0; i < OWN; ++i)
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    if (i >= R.nown) continue;                     // (warp-uniform)
                    const float4 a = add4[i][e];
                    ag[i][e] = fmaf(acc[i][4 * e + 2], scale, a.z);
                    e1[i][e] = tc_exp(-tc_clamp(fmaf(acc[i][4 * e], scale, a.x)));
                    e2[i][e] = tc_exp(-tc_clamp(fmaf(acc[i][4 * e + 1], scale, a.y)));
                    e3[i][e] = tc_exp(-2.f * tc_clamp(ag[i][e]));
                    e4[i][e] = tc_exp(-tc_clamp(fmaf(acc[i][4 * e + 3], scale, a.w)));
                }
#pragma unroll
            for (int i = 0; i < OWN; ++i)
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    if (i >= R.nown) continue;
                    float num, den;
                    klm_tanh_parts(ag[i][e], e3[i][e], num, den);
                    const float ig = __fdividef(num, (1.f + e1[i][e]) * den);
                    const float f = __fdividef(1.f, 1.f + e2[i][e]);
                    e1[i][e] = fmaf(f, c_reg[i][e], ig);                        // new cell state
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
                    if (active) { c_reg[i][e] = e1[i][e]; h_reg[i][e] = hn; }
                }
        }
        if (step && !active) {
#pragma unroll
            for (int i = 0; i < OWN; ++i) h_reg[i][0] = h_reg[i][1] = 0.f;
        }
        stage();
        if (TRACE && tid == 64) tc_trace(p, tick, 6);
        tc_bar_epi();
        if (warp == 2 && !bp_ok && alive)
            alive = df_wait_all(p, p.sync + KLM_DF_HRD + l * 64, role.n_readers, tick_base + (unsigned)(tick - KLM_DF_R + 1), 9);
        if (tid == 64) {
            const unsigned version = (unsigned)tick + 2u;
            bulk_out(hx_l + (size_t)(version & (KLM_DF_R - 1)) * hx_slice + (size_t)(role.first_pair / 4) * (128 * 8));
            if (TRACE && (p.exp & 128)) tc_trace(p, tick, 6);          // (tc_exp bit 128: 'staged' = bulk store performed)
            // everything this CTA hands over went through that one bulk store and it has been performed (wait_group 0): a
            // release has nothing left to order (klm_set_option "tc_pub_release" restores it)
            if (p.fault && role.rot == 0 && tick == 2) { }                       // (test hook: the watchdog must end the launch)
            else df_signal(my_flag, epoch + version, p.pub_release);
            if (l > 0) df_st_relaxed(p.sync + KLM_DF_ZRD + l * 64 + role.idx, tick_base + (unsigned)tick + 1u);   // Z of this tick consumed (all threads: bar.sync above)
            if (x.guard) tc::mbar_arrive(x.bar_epid);
            if (TRACE) tc_trace(p, tick, 7);
        }
        if (step && active) {
            // step mode: the new state goes straight to the pool (the hypothesis' own output slot) -- BEHIND the publication:
            // nobody reads it in this launch, and in front of it the 32-sector stores sat in the memory pipeline before the
            // staging stores and under the release of the hand-over (~1 us per hop of the beam step's chain)
            float *st = p.state + ((size_t)slot_out * p.D + l) * 2 * W;
#pragma unroll
            for (int i = 0; i < OWN; ++i)
                if (i < R.nown) {
                    const int u = 2 * (pr0 + i);
                    *reinterpret_cast<float2 *>(st + u) = make_float2(h_reg[i][0], h_reg[i][1]);
                    *reinterpret_cast<float2 *>(st + W + u) = make_float2(c_reg[i][0], c_reg[i][1]);
                }
        }
        alive = __all_sync(0xffffffffu, alive ? 1 : 0) != 0;
    }
    // ---- sequence mode: the final state goes back to the pool ----
    if (!step && s < p.B) {
        float *st = p.state + ((size_t)p.slots_out[s] * p.D + l) * 2 * W;
#pragma unroll
        for (int i = 0; i < OWN; ++i)
            if (i < R.nown) {
                const int u = 2 * (pr0 + i);
                *reinterpret_cast<float2 *>(st + u) = make_float2(h_reg[i][0], h_reg[i][1]);
                *reinterpret_cast<float2 *>(st + W + u) = make_float2(c_reg[i][0], c_reg[i][1]);
            }
    }
}

// Epilogue of an IN / PROJ CTA: the hi rows' part (hi.hi + hi.lo, + bias) and the lo rows' part (lo.hi) of the slice leave
// as two arrays, the consumer adds them: no hand-over between threads, nothing aliased on the ring.
template <bool TRACE>
__device__ __noinline__ void df_epi_inproj(const KlmTcParams &p, const KlmTcRole &role, const DfCtx &x) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int W = p.W, NP = x.NP, N = x.N, T = x.T, l = role.layer;
    const int q = warp & 3, row = 32 * q + lane, half = row >> 6, s = row & 63, ch = (warp - 2) >> 2;
    const bool is_in = role.kind == KLM_TC_IN;
    const int k_first = ch ? (NP + 1) / 2 : 0, k_end = ch ? NP : (NP + 1) / 2;      // pairs of this warp
    const uint32_t trow = x.tmem_base + ((uint32_t)(32 * q) << 16);
    const float scale = role.scale;
    const unsigned tick_base = x.tick_base;
    unsigned *done_word = p.sync + (is_in ? KLM_DF_ZPROG + l * 64 : KLM_DF_PPROG) + role.idx;
    unsigned *read_word = p.sync + KLM_DF_HRD + role.src_layer * 64 + role.idx;
    const int nslots = is_in ? KLM_DF_NZ : KLM_DF_NL;
    const bool deferred = x.guard != 0;
    if (tid == 64 && deferred) tc::mbar_arrive(x.bar_epid);          // (arrival 0: the producer may fetch tick 0)
    bool alive = true;
    for (int tick = 0; tick < T; ++tick) {
        // back-pressure: the slot written now was last read in tick - nslots (IN: by the REC CTAs of the layer, a counter;
        // PROJ: by the LSE warps, one progress word each)
        if (tick >= nslots && alive) {
            if (is_in) {
                alive = df_wait_all(p, p.sync + KLM_DF_ZRD + l * 64, role.n_zcons, tick_base + (unsigned)(tick - nslots + 1), 11);
            } else {
                // LSE warp w works on ticks w, w + n, w + 2n, ...: its word holds tick_base + (last finished tick + 1)
                const unsigned need_tick = (unsigned)(tick - nslots);          // must be finished by whoever owns it
                const int owner = (int)(need_tick % (unsigned)p.n_lse_warps);
                int ok = 1;
                if (lane == 0) ok = df_wait_word(p, p.sync + KLM_DF_LPROG + owner, tick_base + need_tick + 1u, 12) ? 1 : 0;
                alive = __all_sync(0xffffffffu, ok) != 0;
            }
        }
        {
            int ok = alive ? (tc_wait_mbar(x.bar_tfull, tick & 1, p, 6) ? 1 : 0) : 1;
            alive = alive && __all_sync(0xffffffffu, ok);
        }
        if (tid == 64) {
            df_st_relaxed(read_word, tick_base + (unsigned)tick + 1u);   // every MMA of the tick is complete: the operand has been read
            if (TRACE) tc_trace(p, tick, 4);
            if (deferred) {
                // the quiet moment of this CTA: every bulk copy of the tick has landed, the producer holds the next ones back.
                // A gpu-scope release issued while copies are streaming in takes ~4 600 clk instead of ~900 (measured), so
                // what the previous tick stored is published HERE, one tick late, and only then may the producer go on
                if (tick > 0) df_st_release(done_word, tick_base + (unsigned)tick);
                tc::mbar_arrive(x.bar_epid);
                if (TRACE && tick > 0) tc_trace(p, tick - 1, 7);
            }
        }
        tc::tc_fence_after();
        float4 *dst;
        if (is_in) dst = reinterpret_cast<float4 *>(p.Z) + ((size_t)(l * KLM_DF_NZ + (tick & (KLM_DF_NZ - 1))) * 2 + half) * W * 64 + s;
        else dst = reinterpret_cast<float4 *>(p.logits) + ((size_t)(tick & (KLM_DF_NL - 1)) * 2 + half) * (p.Vp / 4) * 64 + s;
#pragma unroll 2
        for (int k = k_first; k < k_end; ++k) {
            float a[8];
            tc::tmem_ld8(trow + 8 * k, a);
            if (half == 0) {
                float b[8];
                tc::tmem_ld8(trow + N + 8 * k, b);
                tc::tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 8; ++j) a[j] += b[j];
            } else {
                tc::tmem_ld_wait();
            }
            const int u = 2 * (role.first_pair + k);                // IN: first unit of the pair; PROJ: first float4 of 8 vocabulary rows
            float4 b0 = make_float4(0.f, 0.f, 0.f, 0.f), b1 = b0;
            if (is_in && half == 0) { b0 = __ldg(p.bias[l] + u); b1 = __ldg(p.bias[l] + u + 1); }
            __stcg(dst + (size_t)u * 64, make_float4(fmaf(a[0], scale, b0.x), fmaf(a[1], scale, b0.y), fmaf(a[2], scale, b0.z), fmaf(a[3], scale, b0.w)));
            __stcg(dst + (size_t)(u + 1) * 64, make_float4(fmaf(a[4], scale, b1.x), fmaf(a[5], scale, b1.y), fmaf(a[6], scale, b1.z), fmaf(a[7], scale, b1.w)));
        }
        tc::tc_fence_before();
        tc::mbar_arrive(x.bar_tempty);
        if (TRACE && tid == 64) tc_trace(p, tick, 5);
        tc_bar_epi();                                                // all stores of the CTA issued ...
        if (tid == 64 && !deferred) {
            df_st_release(done_word, tick_base + (unsigned)tick + 1u);   // ... and published (release: cumulative over the bar.sync)
            if (TRACE) tc_trace(p, tick, 7);
        }
    }
    if (tid == 64 && deferred) df_st_release(done_word, tick_base + (unsigned)T);
}

// ---------------------------------------------------------------------------------------------
// LSE stage: every warp on its own -- lane = stream (32 streams), the whole vocabulary, ticks w, w + n, w + 2n, ...
// Adds the two parts of the logits, log-sum-exp in one pass, picks the target's logit (or writes the softmax row).
__device__ void df_lse_role(const KlmTcParams &p, const KlmTcRole &role, const DfCtx &x) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int V = p.V, Vq = p.Vp / 4, V4 = (V + 3) / 4, full_rows = p.full_rows;
    const int T = x.groups > 0 ? x.groups : p.T_max;
    const bool step = x.groups > 0;
    const int ngroups = (step ? 64 : p.B) > 32 ? 2 : 1;            // stream groups of 32 per tick
    const int gw = warp + (KLM_TC_THREADS / 32) * role.first_pair;  // global LSE warp index (several LSE CTAs: interleaved)
    const int nw = p.n_lse_warps;
    if (gw >= nw) return;
    unsigned *my_prog = p.sync + KLM_DF_LPROG + gw;
    const unsigned tick_base = x.tick_base;
    const int *targets = p.targets;
    // work items: (tick, group) pairs, item k = tick * ngroups + group; warp gw takes the ticks gw, gw + nw, ... and all
    // groups of a tick (so that "tick finished" is one word per warp)
    // sequence mode: warp gw takes the ticks gw, gw + nw, ... with all their stream groups ("tick finished" is one word per
    // warp); step mode: the few (tick, group) items are spread over the warps -- a beam step waits for the last of them
    const int nitems = step ? T * ngroups : T;
    for (int item = gw; item < nitems; item += nw) {
        const int tick = step ? item / ngroups : item;
        const int g_first = step ? item % ngroups : 0, g_end = step ? g_first + 1 : ngroups;
        if (!df_wait_all(p, p.sync + KLM_DF_PPROG, p.n_proj, tick_base + (unsigned)tick + 1u, 10)) return;
        for (int g = g_first; g < g_end; ++g) {
            const int s = 32 * g + lane;
            const int hyp = step ? 64 * tick + s : s;
            long long o0 = 0;
            int len = 0;
            if (!step && s < p.B) { o0 = p.offsets[s]; len = (int)(p.offsets[s + 1] - o0); }
            const bool act = step ? hyp < x.n_hyp : (s < p.B && tick < len);
            const long long pos = step ? hyp : o0 + tick;
            const float4 *pa = reinterpret_cast<const float4 *>(p.logits) + ((size_t)(tick & (KLM_DF_NL - 1)) * 2) * Vq * 64 + s;
            const float4 *pb = pa + (size_t)Vq * 64;
            float tl = 0.f;
            if (act && !full_rows) {
                int tg = (step && p.step_inline) ? p.step.targets[hyp] : __ldg(targets + pos);
                tg = min(max(tg, 0), V - 1);
                tl = __ldcg(reinterpret_cast<const float *>(pa + (size_t)(tg >> 2) * 64) + (tg & 3)) +
                     __ldcg(reinterpret_cast<const float *>(pb + (size_t)(tg >> 2) * 64) + (tg & 3));
            }
            float m = -INFINITY, sum = 0.f;
            for (int v0 = 0; v0 < V4; v0 += 8) {          // 16 loads in flight per lane: the loop is pure L2 latency
                float4 xv[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    if (v0 + j < V4) {
                        const float4 a = __ldcg(pa + (size_t)(v0 + j) * 64), b = __ldcg(pb + (size_t)(v0 + j) * 64);
                        xv[j] = make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
                        const int v = 4 * (v0 + j);
                        if (v + 1 >= V) xv[j].y = -INFINITY;
                        if (v + 2 >= V) xv[j].z = -INFINITY;
                        if (v + 3 >= V) xv[j].w = -INFINITY;
                    } else {
                        xv[j] = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
                    }
                }
                float mx = m;
#pragma unroll
                for (int j = 0; j < 8; ++j) mx = fmaxf(fmaxf(fmaxf(mx, xv[j].x), fmaxf(xv[j].y, xv[j].z)), xv[j].w);
                float acc = sum * __expf(m - mx);
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    acc += __expf(xv[j].x - mx) + __expf(xv[j].y - mx) + __expf(xv[j].z - mx) + __expf(xv[j].w - mx);
                sum = acc; m = mx;
            }
            if (act) {
                if (!full_rows) {
                    p.out[pos] = (tl - m) - logf(sum);
                } else {
                    float *dst = p.out + (size_t)pos * V;
                    const float inv = 1.0f / sum;
                    for (int v4 = 0; v4 < V4; ++v4) {
                        const float4 a = __ldcg(pa + (size_t)v4 * 64), b = __ldcg(pb + (size_t)v4 * 64);
                        const int v = 4 * v4;
                        dst[v] = expf(a.x + b.x - m) * inv;
                        if (v + 1 < V) dst[v + 1] = expf(a.y + b.y - m) * inv;
                        if (v + 2 < V) dst[v + 2] = expf(a.z + b.z - m) * inv;
                        if (v + 3 < V) dst[v + 3] = expf(a.w + b.w - m) * inv;
                    }
                }
            }
        }
        __syncwarp();
        if (lane == 0 && !step) df_st_relaxed(my_prog, tick_base + (unsigned)tick + 1u);    // the logits of this tick have been consumed
    }
}

__device__ __forceinline__ unsigned df_ld_acquire_sys(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Resident instance (KlmTcParams::server): waits for step `seq`.  The first LSE CTA polls the host's command block in
// mapped memory, copies what the step needs (header, the arrays of its n_hyp hypotheses) to the device block and rings the
// device-side doorbell; every other CTA polls that.  Fills the per-step fields of x; false = leave the kernel (exit
// command, or the watchdog: nobody rang for spin_limit polls).
__device__ __forceinline__ bool df_next_step(const KlmTcParams &p, const KlmTcRole &role, unsigned seq, DfCtx &x) {
    __shared__ int s_go;
    const int tid = threadIdx.x;
    KlmStepCmd *cd = p.cmd_dev;
    if (role.kind == KLM_TC_LSE && role.first_pair == 0) {
        const KlmStepCmd *ch = p.cmd_host;
        if (tid == 0) {
            unsigned n = 0;
            int ok = 1;
            while (df_ld_acquire_sys(&ch->seq) != seq)
                if (tc_spin_fail(n, p, 14)) { ok = 0; break; }
            s_go = ok;
        }
        __syncthreads();
        if (!s_go) return false;
        unsigned long long t_seen = 0;
        if (tid == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_seen));
        // ONE pass of 16-byte loads over the bus (everything in flight at once: one round trip): the header by thread 0, the
        // four per-hypothesis arrays (4 KB) and the context indices by everybody; entries beyond n_hyp are stale and unused
        {
            const uint4 *src = reinterpret_cast<const uint4 *>(&ch->args);
            uint4 *dst = reinterpret_cast<uint4 *>(&cd->args);
            const int n16 = KLM_STEP_MAX + (KLM_STEP_MAX * p.Nc + 3) / 4;          // uint4 pieces: 4 arrays of 256 ints, then ctx
            uint4 v0 = make_uint4(0, 0, 0, 0), v1 = v0, h0 = v0, h1 = v0;
            if (tid < n16) v0 = src[tid];
            if (tid + KLM_TC_THREADS < n16) v1 = src[tid + KLM_TC_THREADS];
            if (tid == 0) { h0 = reinterpret_cast<const uint4 *>(ch)[0]; h1 = reinterpret_cast<const uint4 *>(ch)[1]; }
            if (tid < n16) dst[tid] = v0;
            if (tid + KLM_TC_THREADS < n16) dst[tid + KLM_TC_THREADS] = v1;
            if (tid == 0) { cd->n_hyp = (int)h0.y; cd->groups = (int)h0.z; cd->epoch = h0.w; cd->tick_base = h1.x; }
        }
        __threadfence();
        __syncthreads();
        if (tid == 0) {
            df_st_release(&cd->seq, seq);
            // (diagnostics, words 8..11 of the completion block: when the doorbell was seen / the device-side bell rung, ns)
            unsigned long long t_rung;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_rung));
            *(volatile unsigned long long *)(p.srv_done + 8) = t_seen;
            *(volatile unsigned long long *)(p.srv_done + 10) = t_rung;
        }
    } else {
        if (tid == 0) {
            unsigned n = 0;
            int ok = 1;
            while (!df_reached(df_ld_acquire(&cd->seq), seq)) {
                if (tc_spin_fail(n, p, 15)) { ok = 0; break; }
                __nanosleep(100);
            }
            s_go = ok;
        }
        __syncthreads();
        if (!s_go) return false;
    }
    x.n_hyp = *(volatile int *)&cd->n_hyp; x.groups = *(volatile int *)&cd->groups;
    x.epoch = *(volatile unsigned *)&cd->epoch; x.tick_base = *(volatile unsigned *)&cd->tick_base;
    x.T = x.groups;
    __syncthreads();                                    // (s_go is reused by the next step)
    return x.n_hyp > 0;
}

template <bool TRACE>
__device__ __forceinline__ void df_body(const KlmTcParams &p, unsigned char *smem) {
    if (TRACE && threadIdx.x == 64) tc_trace(p, p.trace_t0 + KLM_TC_TRACE_TICKS - 1, 0);      // (last trace slot: phases of the launch)
    const KlmTcRole role = p.roles[blockIdx.x];
    const int tid = threadIdx.x, warp = tid >> 5;
    DfCtx x;
    x.epoch = p.epoch; x.tick_base = p.tick_base; x.n_hyp = p.n_hyp; x.groups = p.step_groups;
    x.T = p.step_groups > 0 ? p.step_groups : p.T_max;
    if (role.kind == KLM_TC_LSE) {
        x.ring_a = x.bar_full = x.bar_empty = x.bar_tfull = x.bar_tempty = x.bar_epid = x.tmem_base = x.wsm_a = x.masks_a = 0;
        x.N = x.NP = x.crot = x.guard = x.nch = 0; x.W = p.W;
        for (unsigned seq = 1;; ++seq) {
            if (p.server && !df_next_step(p, role, seq, x)) break;
            df_lse_role(p, role, x);
            __syncthreads();
            if (!p.server) break;
            // the scores of this CTA's share are in the host's buffer: completion word (mapped host memory)
            if (tid == 0) {
                if (role.first_pair == 0) {
                    unsigned long long t_done;
                    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_done));
                    *(volatile unsigned long long *)(p.srv_done + 12) = t_done;
                }
                // a wait of this step that timed out anywhere (this CTA's own included) is reported with the completion
                const int code = *(volatile int *)p.err;
                if (code && p.err_host) *(volatile int *)p.err_host = code;
                __threadfence_system();
                *(volatile unsigned *)(p.srv_done + role.first_pair) = seq;
            }
        }
        __syncthreads();
        if (tid == 0 && p.err_host) {
            const int code = *(volatile int *)p.err;
            if (code) *(volatile int *)p.err_host = code;
        }
        return;
    }

    const int W = p.W, S = role.stages, kc = role.kc;
    const int NP = role.npairs, N = 8 * NP;

    // ---- shared memory: [ring of S chunks (REC: hand-over + staging alias one slot)][barriers + chunk masks][B image] ----
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + role.ring_bytes);
    unsigned char *wsm = smem + role.ring_bytes + 256;
    const uint32_t bar_full = tc::smem_u32(bars), bar_empty = tc::smem_u32(bars + 4);
    const uint32_t bar_tfull = tc::smem_u32(bars + 8), bar_tempty = tc::smem_u32(bars + 9);
    const uint32_t bar_epid = tc::smem_u32(bars + 10);
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 11);
    unsigned long long *masks = reinterpret_cast<unsigned long long *>(bars + 16);        // [16]
    if (warp == 0) {
        tc::tmem_alloc(tc::smem_u32(tmem_slot), KLM_TC_TMEM_COLS);
        tc::tmem_relinquish();
    }
    auto init_bars = [&]() {         // (thread 32)
        for (int i = 0; i < S; ++i) { tc::mbar_init(bar_full + 8 * i, 1); tc::mbar_init(bar_empty + 8 * i, 1); }
        tc::mbar_init(bar_tfull, 1);
        tc::mbar_init(bar_tempty, 256);
        tc::mbar_init(bar_epid, 1);
    };
    if (tid == 32) {
        init_bars();
        tc::mbar_init(tc::smem_u32(bars + 13), 1);       // B image landed
        tc::fence_mbar_init();
    }
    if (tid >= 64 && tid < 80) masks[tid - 64] = role.ch_mask[tid - 64];
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    // B image: global -> shared, once per launch, as bulk async copies on a barrier of their own (the MMA warp waits for it
    // before its first MMA; everything else of the prologue overlaps the ~130-180 KB transfer)
    const uint32_t bar_img = tc::smem_u32(bars + 13);
    if (tid == 32) {
        const uint32_t img_bytes = (uint32_t)W * 2u * (uint32_t)N * 2u;
        tc::mbar_arrive_expect_tx(bar_img, img_bytes);
        for (uint32_t off = 0; off < img_bytes; off += 32768u)
            tc::bulk_g2s(tc::smem_u32(wsm) + off, p.wimg + role.wofs + off, min(32768u, img_bytes - off), bar_img);
    }

    if (TRACE && tid == 64) tc_trace(p, p.trace_t0 + KLM_TC_TRACE_TICKS - 1, 1);
    x.ring_a = tc::smem_u32(smem); x.bar_full = bar_full; x.bar_empty = bar_empty; x.bar_tfull = bar_tfull;
    x.bar_tempty = bar_tempty; x.bar_epid = bar_epid; x.tmem_base = *tmem_slot; x.wsm_a = tc::smem_u32(wsm);
    x.masks_a = tc::smem_u32(masks);
    x.N = N; x.NP = NP; x.W = W; x.nch = W / kc;
    const bool stepm = p.step_groups > 0 || p.server;
    // REC: the producer waits for the epilogue where the aliased buffers are not covered by the own-flag rule;
    // IN / PROJ: the producer waits until the epilogue has published the previous tick (at the quiet moment after the MMAs)
    // (step mode: IN / PROJ publish at once -- a beam step is about latency, and there is little to stream in)
    x.guard = role.kind == KLM_TC_REC ? ((role.guard || stepm || (p.exp & 2)) ? 1 : 0)
                                      : (((p.exp & 4) || stepm) ? 0 : 1);
    // REC: the chunk with the CTA's own slice goes into ring slot `alias_slot`; the others in rotated order behind / before it
    x.crot = role.kind == KLM_TC_REC ? role.sigma : role.rot % x.nch;

    for (unsigned seq = 1;; ++seq) {
        if (p.server) {
            if (!df_next_step(p, role, seq, x)) break;
            if (seq > 1) {
                // the ring / accumulator barriers start every step in their initial phase (the role loops count from 0)
                if (tid == 32) {
                    for (int i = 0; i < S; ++i) {
                        asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(bar_full + 8 * i) : "memory");
                        asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(bar_empty + 8 * i) : "memory");
                    }
                    asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(bar_tfull) : "memory");
                    asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(bar_tempty) : "memory");
                    asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(bar_epid) : "memory");
                    init_bars();
                    tc::fence_mbar_init();
                }
                tc::tc_fence_before();
                __syncthreads();
                tc::tc_fence_after();
            }
        }
        if (warp == 0) {
            if (!df_run_loops<TRACE>(true, p, role, x) && tid == 0) atomicExch(p.err, 90);
        } else if (warp == 1) {
            if (__all_sync(0xffffffffu, tc_wait_mbar(bar_img, 0, p, 13) ? 1 : 0)) df_run_loops<TRACE>(false, p, role, x);
        } else if (role.kind == KLM_TC_REC) {
            unsigned char *alias = smem + (size_t)role.alias_slot * kc * 256;
            const int own = ((NP + 1) / 2 + 1) / 2;
            if (own <= 1) df_epi_rec<1, TRACE>(p, role, x, alias);
            else if (own == 2) df_epi_rec<2, TRACE>(p, role, x, alias);
            else df_epi_rec<3, TRACE>(p, role, x, alias);
        } else {
            df_epi_inproj<TRACE>(p, role, x);
        }
        tc::tc_fence_before();
        __syncthreads();
        tc::tc_fence_after();
        if (!p.server) break;
    }

    if (TRACE && tid == 64) tc_trace(p, p.trace_t0 + KLM_TC_TRACE_TICKS - 1, 3);
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        __syncwarp();
        tc::tmem_dealloc(x.tmem_base, KLM_TC_TMEM_COLS);
    }
    if (TRACE && tid == 64) tc_trace(p, p.trace_t0 + KLM_TC_TRACE_TICKS - 1, 4);
    if (tid == 0 && p.err_host) {
        const int code = *(volatile int *)p.err;
        if (code) *(volatile int *)p.err_host = code;
    }
}

}  // namespace

__global__ void __launch_bounds__(KLM_TC_THREADS, 1) klm_seq_df(const __grid_constant__ KlmTcParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    df_body<false>(p, smem);
}
__global__ void __launch_bounds__(KLM_TC_THREADS, 1) klm_seq_df_traced(const __grid_constant__ KlmTcParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    df_body<true>(p, smem);
}

cudaError_t klm_launch_seq_df(const KlmTcParams &p, int grid, size_t smem, cudaStream_t stream) {
    void *args[] = { const_cast<KlmTcParams *>(&p) };
    static size_t smem_set[2][64] = {{0}};
    const int which = p.trace ? 1 : 0;
    const void *fn = which ? (const void *)klm_seq_df_traced : (const void *)klm_seq_df;
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64 || smem_set[which][dev] != smem) {
        cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        if (dev >= 0 && dev < 64) smem_set[which][dev] = smem;
    }
    // All CTAs must be co-resident (they wait for one another).  The grid never exceeds one CTA per SM and the model's
    // stream runs one kernel at a time, so a plain launch has them resident as well -- and costs ~25 us less per launch than
    // a cooperative one, which matters for beam steps (a device-side watchdog turns a violation into an error, not a hang).
    // klm_set_option "tc_exp" bit 8 asks for the cooperative launch.
    if (p.exp & 8) return cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(KLM_TC_THREADS), args, smem, stream);
    return cudaLaunchKernel(fn, dim3(grid), dim3(KLM_TC_THREADS), args, smem, stream);
}
