// One text at a time (Rater.rate / rate2 of a single string -- the reference's default use, rating.py:493-576): the
// whole model in the REGISTERS of one thread-block cluster, the recurrence exchanged through distributed shared
// memory.  Replaces the per-window Session.run of Rater.rate (rating.py:516) for ONE stream and small models
// (the reference's default: width 128, depth 2).
//
// With one stream a tick is pure latency.  klm_seq_fma_ll (weights in shared memory, {value, tick} words through
// L2) needs ~4 900 clk per tick, of which ~2 500 are the L2 round trips and most of the rest is one warp walking a
// whole row of weights (8 k-steps of two shared-memory loads, 10 shuffles); the first cluster kernel (klm_seq_cl,
// round 1, now scripts/ubench/) removed the L2 trips but kept the walk: 4.2 us.  Here
//   * CTA c of a cluster of C = 16 (8) owns units [8 c, 8 c + 8) of EVERY layer and V / C rows of the tied output
//     projection; warp w owns unit 8 c + w of every layer: a warp IS one LSTM unit, lane = k-segment (values
//     4 lane .. 4 lane + 3 of every input vector), and the lane keeps the weights of ALL FOUR gates for its four values
//     in registers for the whole launch (width 128, depth 2, 256 characters: 16 + 32 + 8 = 56 registers);
//   * per tick a lane loads ONE float4 per hidden vector (a warp reads every value once: 3 x 512 B per tick; with
//     lane = (gate, k-segment) the four gate groups each loaded the whole vector and the eight unit warps together kept
//     the shared-memory pipe busy for ~350 clk of every tick), does 4 (8) FMAs per gate, and the 4 D gate sums are
//     reduced over the 32 lanes by recursive halving (a lane sends half of its values and keeps the other half:
//     4 + 2 + 1 + 1 + 1 shuffles for depth 2), four indexed shuffles bring the gates of layer l to lane l, lanes
//     0..D-1 run the cells side by side (8 MUFU operations) and lanes 0..C-1 store the new values of ALL layers as
//     adjacent 8-byte {value, tick} words with ONE vector store into the ring of every CTA of the cluster
//     (st.shared::cluster, ~215 clk) -- no fence, no flag, no barrier across CTAs.  (One store per layer cost 380 clk
//     of every tick: the 16-destination stores are what the publication pays for.);
//   * one more warp per CTA polls the CTA's own ring, copies a complete hidden vector into a dense array and
//     releases the unit warps through a named barrier (sixteen warps spinning on shared memory starve
//     the port the remote stores arrive through); in CTA 0 another warp does the log-sum-exp + target pick of the
//     position that left the projection a tick earlier.
// Wavefront as in the other kernels: layer l works on position tick - l, the projection on tick - D, the
// log-sum-exp on tick - D - 1; every CTA takes part in layer 1's recurrence, so nobody is more than a tick ahead of
// anybody and a ring of 4 ticks needs no back-pressure except for the log-sum-exp (one progress word in CTA 0).
#include <cstdio>
#include <cstring>
#include "klm_common.cuh"

// CTAs per cluster: 16 (non-portable size; 8 units per CTA, 9 warps, no register pressure) where the device allows it,
// else 8 (16 units per CTA, 17 warps: 96 registers per thread, a few weights spill)
#define RG_RING      4
#define RG_SPIN      (1u << 22)

namespace {

__device__ __forceinline__ uint32_t rg_smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t rg_mapa(uint32_t local, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
    return r;
}
__device__ __forceinline__ void rg_st_remote(uint32_t addr, float v, unsigned tick) {
    asm volatile("st.shared::cluster.v2.u32 [%0], {%1, %2};" ::"r"(addr), "r"(__float_as_uint(v)), "r"(tick) : "memory");
}
__device__ __forceinline__ uint2 rg_ld_word(const uint2 *p) {
    uint2 v;
    asm volatile("ld.volatile.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(rg_smem_u32(p)) : "memory");
    return v;
}
__device__ __forceinline__ void rg_cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
template <int NT> __device__ __forceinline__ void rg_bar_all() { asm volatile("bar.sync 1, %0;" ::"n"(NT) : "memory"); }
__device__ __forceinline__ float rg_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x * 1.4426950408889634f));
    return y;
}
__device__ __forceinline__ void add4(float4 &a, const float4 b) { a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }
__device__ __forceinline__ float rg_clamp(float x) { return fminf(fmaxf(x, -30.f), 30.f); }
// LSTM cell (Keras LSTMCell, gate order i|f|c|o, exact logistic: rating.py:133-134); 8 MUFU operations
__device__ __forceinline__ void rg_cell(float ai, float af, float ag, float ao, float &c, float &h) {
    const float ei = rg_ex2(-rg_clamp(ai)), ef = rg_ex2(-rg_clamp(af)), eg = rg_ex2(-2.f * rg_clamp(ag)), eo = rg_ex2(-rg_clamp(ao));
    float num, den;
    klm_tanh_parts(ag, eg, num, den);
    const float ig = __fdividef(num, (1.f + ei) * den);
    const float f = __fdividef(1.f, 1.f + ef);
    c = fmaf(f, c, ig);
    const float ec = rg_ex2(-2.f * rg_clamp(c));
    klm_tanh_parts(c, ec, num, den);
    h = __fdividef(num, (1.f + eo) * den);
}

}  // namespace

// WS = W / 8 (values per k-segment), D = depth, PR = rows of the output projection per warp (V <= 128 * PR)
template <int WS, int D, int PR, int C>
__global__ void __launch_bounds__(32 * (8 * WS / C + 2), 1) klm_seq_rg(const __grid_constant__ KlmSeqParams p) {
    constexpr int W = 8 * WS, R = RG_RING;
    constexpr int UPC = W / C;                   // units of every layer per CTA = unit warps
    constexpr int RG_THREADS = 32 * (UPC + 2);   // unit warps, the poll warp, the log-sum-exp warp (at work in CTA 0 only)
    constexpr int BAR_THREADS = 32 * (UPC + 1);  // the per-tick barrier: unit warps + poll warp
    constexpr int LPR = 32 / PR;                 // lanes per projection row
    constexpr int PS = W / LPR;                  // values of the hidden vector per lane in the projection
    constexpr int VP = W * PR;                   // projection rows of the cluster (>= V)
    constexpr int DP = D == 1 ? 1 : (D + 1) / 2 * 2;   // words per unit in a ring slot: the D layers side by side, padded to 16 bytes
    constexpr int HW = W * DP;                   // hidden words of one ring slot
    constexpr int SLOT = HW + VP;                // words of one ring slot
    constexpr int NVP = D == 1 ? 4 : D == 2 ? 8 : 16;  // gate sums per lane (4 D), padded to a power of two for the halving
    constexpr int LPV = 32 / NVP;                // lanes that end up with the same gate sum
    static_assert(W == 128, "a lane owns four values of a 128-wide vector");
    extern __shared__ __align__(16) unsigned char rg_smem[];
    uint2 *ring = reinterpret_cast<uint2 *>(rg_smem);
    float *dense = reinterpret_cast<float *>(ring + (size_t)R * SLOT);          // [2][D][W]
    float4 *tab = reinterpret_cast<float4 *>(dense + 2 * D * W);               // [16 warps][D + 1] tick-invariant addends (lane 0 only:
                                                                                 //  kept out of the registers, which hold the weights)
    int *lse_done = reinterpret_cast<int *>(tab + UPC * (D + 1));      // CTA 0: ticks the log-sum-exp warp is through with
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned rank;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    const int nticks = p.T_max + D + 1;
    const long long o0 = p.offsets[0];
    const int len = (int)(p.offsets[1] - o0);

    for (int i = threadIdx.x; i < R * SLOT; i += RG_THREADS) ring[i] = make_uint2(0u, 0xFFFFFFFFu);
    if (threadIdx.x == 0) *lse_done = D + 1;         // (no logits before tick D + 1)
    __syncthreads();
    rg_cluster_sync();                            // every ring is initialised before anybody stores into it

    if (warp < UPC) {
        // =========================== unit warp: unit u of every layer, PR rows of the projection ===========================
        const int u = (int)rank * UPC + warp;
        // ---- weights -> registers: values k = 4 lane + e of every input vector, the four gates of unit u as one float4 ----
        float4 wr[D][4], wi[D > 1 ? D - 1 : 1][4];
        float wp[PS];
#pragma unroll
        for (int l = 0; l < D; ++l) {
            const int K = l > 0 ? 2 * W : W;
            const float4 *src = p.Wl[l] + (size_t)u * K + 4 * lane;                                  // float4 per k: (i, f, g, o)
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                if (l > 0) wi[l > 0 ? l - 1 : 0][e] = __ldg(src + e);                                 // K_l rows (input h_{l-1})
                wr[l][e] = __ldg(src + (l > 0 ? W : 0) + e);                                          // R_l rows (input h_l)
            }
        }
        // projection row v = u PR + prow: lane (prow, pseg) takes the float4s pseg + LPR e of the last hidden vector
        const int prow = lane / LPR, pseg = lane % LPR;
        const int v = u * PR + prow;
#pragma unroll
        for (int j = 0; j < PS; ++j) {
            const int k = 4 * (pseg + LPR * (j >> 2)) + (j & 3);
            wp[j] = v < p.V ? __ldg(reinterpret_cast<const float *>(p.Ep + (size_t)(v >> 2) * W + k) + (v & 3)) : 0.f;
        }
        // ---- state, tick-invariant addends (lane 0 runs the cells) ----
        float c_own = 0.f, h_own = 0.f;                    // lane l < D: cell and hidden state of layer l
        float4 *mytab = tab + warp * (D + 1);              // [0] context part (real positions), [l] bias of layer l, [D] context 0
        if (lane < D) {
            const int slot = p.slots_in[0];
            if (slot >= 0) {
                const float *st = p.state + ((size_t)slot * D + lane) * 2 * W;
                h_own = st[u]; c_own = st[W + u];
            }
            if (lane > 0) mytab[lane] = __ldg(p.bias[lane] + u);
        }
        if (lane == 0) {
            float4 qreal = make_float4(0.f, 0.f, 0.f, 0.f), qpad = qreal;
            for (int k = 0; k < p.Nc; ++k) {
                int z = __ldg(p.ctx + k);
                z = min(max(z, 0), p.ctx_rows[k] - 1);
                add4(qreal, __ldg(p.Q[k] + (size_t)z * W + u));
                add4(qpad, __ldg(p.Q[k] + u));
            }
            mytab[0] = qreal; mytab[D] = qpad;
        }
        __syncwarp();
        const int nreal = p.n_real ? __ldg(p.n_real) : 0x7fffffff;
        // remote addresses of this warp's words in ring slot 0 of CTA `lane` (lanes 0..C-1), and of its logits in CTA 0
        const uint32_t dst_h0 = rg_mapa(rg_smem_u32(ring + u * DP), (uint32_t)(lane & (C - 1)));
        const uint32_t dst_p = rg_mapa(rg_smem_u32(ring + HW + v), 0u);
        // the carried hidden state is the word of tick 0
        {
            float h0[D];
#pragma unroll
            for (int l = 0; l < D; ++l) h0[l] = __shfl_sync(0xffffffffu, h_own, l);
            if (lane < C) {
#pragma unroll
                for (int l = 0; l < D; ++l) rg_st_remote(dst_h0 + 8u * l, h0[l], 0u);
            }
        }
        // layer 1's table rows travel a tick ahead of their use (the text is known): row of position 0 now
        // (pre_cur = row of position tick, pre_nxt = row of tick + 1, x_nn = character of tick + 2, x_n2 = of tick + 3)
        int x_nn = 0, x_n2 = 0;
        float4 pre_cur = make_float4(0.f, 0.f, 0.f, 0.f), pre_nxt = pre_cur;
        if (lane == 0 && len > 0) {
            pre_cur = __ldg(p.P + (size_t)min(max(__ldg(p.chars + o0), 0), p.V - 1) * W + u);
            if (len > 1) pre_nxt = __ldg(p.P + (size_t)min(max(__ldg(p.chars + o0 + 1), 0), p.V - 1) * W + u);
            if (len > 2) x_nn = __ldg(p.chars + o0 + 2);
            if (len > 3) x_n2 = __ldg(p.chars + o0 + 3);
        }

        long long dbg[5] = {0, 0, 0, 0, 0}, dbg5 = 0, dbg_t = clock64();
#define RG_DBG(i) if (p.debug) { const long long now_ = clock64(); dbg[i] += now_ - dbg_t; dbg_t = now_; }
        for (int tick = 0; tick < nticks; ++tick) {
            RG_DBG(0)
            rg_bar_all<BAR_THREADS>();                                   // the poll warp has the hidden vectors of this tick in `dense`
            RG_DBG(1)
            const float4 *hv = reinterpret_cast<const float4 *>(dense + (size_t)(tick & 1) * D * W);
            // ---- contractions of all layers and of the projection (independent: their chains overlap) ----
            float4 h4[D];
#pragma unroll
            for (int l = 0; l < D; ++l) h4[l] = hv[l * (W / 4) + lane];
            if (p.debug && __float_as_uint(h4[0].x) != 0x7fc54321u) dbg5 += clock64() - dbg_t;      // (a clock read that cannot precede the first load's data)
            float sv[NVP];                                               // [4 l + gate]
#pragma unroll
            for (int l = 0; l < D; ++l) {
                float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
#define RG_FMA4(w_, h_) a.x = fmaf((w_).x, (h_), a.x); a.y = fmaf((w_).y, (h_), a.y); a.z = fmaf((w_).z, (h_), a.z); a.w = fmaf((w_).w, (h_), a.w);
                RG_FMA4(wr[l][0], h4[l].x) RG_FMA4(wr[l][1], h4[l].y) RG_FMA4(wr[l][2], h4[l].z) RG_FMA4(wr[l][3], h4[l].w)
                if (l > 0) {
                    const float4 hi = h4[l > 0 ? l - 1 : 0];
                    const float4 *w = wi[l > 0 ? l - 1 : 0];
                    RG_FMA4(w[0], hi.x) RG_FMA4(w[1], hi.y) RG_FMA4(w[2], hi.z) RG_FMA4(w[3], hi.w)
                }
#undef RG_FMA4
                sv[4 * l] = a.x; sv[4 * l + 1] = a.y; sv[4 * l + 2] = a.z; sv[4 * l + 3] = a.w;
            }
#pragma unroll
            for (int j = 4 * D; j < NVP; ++j) sv[j] = 0.f;
            float accp;
            {
                float a0 = 0.f, a1 = 0.f;
#pragma unroll
                for (int e = 0; e < PR; ++e) {
                    const float4 hp = hv[(D - 1) * (W / 4) + pseg + LPR * e];
                    a0 = fmaf(wp[4 * e], hp.x, a0); a1 = fmaf(wp[4 * e + 1], hp.y, a1);
                    a0 = fmaf(wp[4 * e + 2], hp.z, a0); a1 = fmaf(wp[4 * e + 3], hp.w, a1);
                }
                accp = a0 + a1;
            }
            RG_DBG(2)
            // ---- sums over the 32 k-segments by recursive halving: with mask m a lane hands the half of its values that its
            //      partner keeps to lane ^ m; afterwards gate sum j lives (complete) in the lanes j LPV .. j LPV + LPV - 1 ----
#pragma unroll
            for (int half = NVP / 2, m = 16; half >= 1; half >>= 1, m >>= 1) {
                const bool up = (lane & m) != 0;
#pragma unroll
                for (int q = 0; q < half; ++q) {
                    const float send = up ? sv[q] : sv[q + half];
                    const float keep = up ? sv[q + half] : sv[q];
                    sv[q] = keep + __shfl_xor_sync(0xffffffffu, send, m);
                }
            }
#pragma unroll
            for (int m = LPV / 2; m >= 1; m >>= 1) sv[0] += __shfl_xor_sync(0xffffffffu, sv[0], m);
#pragma unroll
            for (int o = 1; o < LPR; o <<= 1) accp += __shfl_xor_sync(0xffffffffu, accp, o);
            // the four gate sums of layer l meet in lane l: the cells of all layers run side by side (lanes 0 .. D-1)
            const int lsrc = (lane < D ? 4 * lane : 0) * LPV;
            const float gi = __shfl_sync(0xffffffffu, sv[0], lsrc), gf = __shfl_sync(0xffffffffu, sv[0], lsrc + LPV),
                        gg = __shfl_sync(0xffffffffu, sv[0], lsrc + 2 * LPV), go = __shfl_sync(0xffffffffu, sv[0], lsrc + 3 * LPV);
            {
                const int t = tick - lane;                                   // lane l: layer l, position tick - l
                if (lane < D && t >= 0 && t < len) {
                    float4 a = lane == 0 ? pre_cur : mytab[lane];
                    if (lane == 0) add4(a, mytab[t < nreal ? 0 : D]);        // padding position: context 0 (:1141-1143)
                    rg_cell(gi + a.x, gf + a.y, gg + a.z, go + a.w, c_own, h_own);
                }
            }
            float hn[D];
#pragma unroll
            for (int l = 0; l < D; ++l) hn[l] = __shfl_sync(0xffffffffu, h_own, l);
            RG_DBG(3)
            // ---- publication: the word of tick + 1 into the ring of every CTA (lanes 0..C-1), logits into CTA 0's ----
            const uint32_t sofs = (uint32_t)((tick + 1) % R) * (uint32_t)SLOT * 8u;
            if (lane < C) {
                if (D >= 2)
                    asm volatile("st.shared::cluster.v4.u32 [%0], {%1, %2, %3, %2};" ::"r"(dst_h0 + sofs), "r"(__float_as_uint(hn[0])), "r"((unsigned)(tick + 1)),
                                 "r"(__float_as_uint(hn[D > 1 ? 1 : 0])) : "memory");
                if (D != 2) rg_st_remote(dst_h0 + (D == 3 ? 16u : 0u) + sofs, hn[D - 1], (unsigned)(tick + 1));
            }
            if (pseg == 0) rg_st_remote(dst_p + sofs, accp, (unsigned)(tick + 1));
            // behind the publication (the warp idles until the next barrier anyway): the table row of the next position and the
            // character index of the one after it
            pre_cur = pre_nxt;
            if (lane == 0) {
                if (tick + 2 < len) pre_nxt = __ldg(p.P + (size_t)min(max(x_nn, 0), p.V - 1) * W + u);
                x_nn = x_n2;
                if (tick + 4 < len) x_n2 = __ldg(p.chars + o0 + tick + 4);
            }
            RG_DBG(4)
        }
        if (p.debug && lane == 0 && (warp == 0 || warp == UPC - 1) && (rank == 0 || rank == C - 1))
            printf("klm_seq_rg cta %u warp %d: per tick: prefetch issue %lld, wait for the poll warp %lld, loads + FMAs %lld (first load's data after %lld), shuffles + cells %lld, publication %lld clk\n",
                   rank, warp, dbg[0] / nticks, dbg[1] / nticks, dbg[2] / nticks, dbg5 / nticks, dbg[3] / nticks, dbg[4] / nticks);
        // ---- final state ----
        if (lane < D) {
            float *st = p.state + ((size_t)p.slots_out[0] * D + lane) * 2 * W;
            st[u] = h_own; st[W + u] = c_own;
        }
    } else if (warp == UPC) {
        // =========================== poll warp ===========================
        bool alive = true;
        constexpr int NPW = HW / 32;                        // ring words per lane and tick (word i: unit i / DP, layer i % DP)
        const bool pad = lane % DP >= D;                    // (32 % DP == 0: a lane sees words of ONE layer, or only padding words)
        const int dofs = (lane % DP) * W + lane / DP;       // word lane + 32 j -> dense[layer][unit]: dofs + j * (32 / DP)
        long long pdbg[3] = {0, 0, 0}, pfirst = 0, pdbg_t = clock64();
#define RG_PDBG(i) if (p.debug) { const long long now_ = clock64(); pdbg[i] += now_ - pdbg_t; pdbg_t = now_; }
        for (int tick = 0; tick < nticks; ++tick) {
            const uint2 *cur = ring + (size_t)(tick % R) * SLOT;
            float *dn = dense + (size_t)(tick & 1) * D * W;
            if (alive) {
                // one pass = all words of this lane loaded side by side; repeated until every one carries this tick's number
                unsigned n = 0;
                bool seen = false;
                for (;;) {
                    uint2 wv[NPW];
                    bool all = true;
#pragma unroll
                    for (int j = 0; j < NPW; ++j) wv[j] = rg_ld_word(cur + lane + 32 * j);
#pragma unroll
                    for (int j = 0; j < NPW; ++j) all = all && wv[j].y == (unsigned)tick;
                    if (p.debug && !seen) {
                        bool anyc = false;
#pragma unroll
                        for (int j = 0; j < NPW; ++j) anyc = anyc || (wv[j].y == (unsigned)tick && !pad);
                        if (__any_sync(0xffffffffu, anyc ? 1 : 0)) { seen = true; pfirst += clock64() - pdbg_t; }
                    }
                    if (__all_sync(0xffffffffu, (all || pad) ? 1 : 0)) {
                        if (!pad) {
#pragma unroll
                            for (int j = 0; j < NPW; ++j) dn[dofs + j * (32 / DP)] = __uint_as_float(wv[j].x);
                        }
                        break;
                    }
                    if (++n > RG_SPIN || ((n & 0x3FFu) == 0 && *(volatile int *)p.err != 0)) { atomicCAS(p.err, 0, 31); alive = false; }
                    alive = __all_sync(0xffffffffu, alive ? 1 : 0) != 0;
                    if (!alive) break;
                }
            }
            RG_PDBG(0)
            // CTA 0: the logits written during tick + 1 land in the slot the log-sum-exp warp read two ticks ago -- it must be
            // through with that (it almost always is: one look at a word in shared memory)
            if (rank == 0 && alive && tick >= 2) {
                unsigned n = 0;
                while (*(volatile int *)lse_done < tick - 2) {
                    if (++n > RG_SPIN) { atomicCAS(p.err, 0, 33); alive = false; break; }
                }
                alive = __all_sync(0xffffffffu, alive ? 1 : 0) != 0;
            }
            __syncwarp();
            RG_PDBG(1)
            rg_bar_all<BAR_THREADS>();
            RG_PDBG(2)
        }
        if (p.debug && lane == 0 && (rank == 0 || rank == C - 1))
            printf("klm_seq_rg cta %u poll warp: per tick: polling + dense copy %lld (first word of the tick after %lld), wait for the log-sum-exp warp %lld, barrier %lld clk\n", rank,
                   pdbg[0] / nticks, pfirst / nticks, pdbg[1] / nticks, pdbg[2] / nticks);
    } else if (rank == 0) {
        // =========================== log-sum-exp warp (CTA 0): position tick - D - 1, whose logits carry this tick's number ===========================
        int tg_next = 0;
        if (!p.full_rows && len > 0) tg_next = __ldg(p.targets + o0);
        bool alive = true;
        long long ldbg[2] = {0, 0}, ldbg_t = clock64();
        for (int tick = D + 1; tick < nticks && alive; ++tick) {
            const int t = tick - D - 1;
            if (t >= len) break;
            const uint2 *lg = ring + (size_t)(tick % R) * SLOT + HW;
            int tg = 0;
            if (!p.full_rows) {
                tg = min(max(tg_next, 0), p.V - 1);
                if (t + 1 < len) tg_next = __ldg(p.targets + o0 + t + 1);
            }
            float x[VP / 32];
            unsigned n = 0;
            for (;;) {                                       // all words of this lane side by side, until every one is there
                uint2 wv[VP / 32];
                bool all = true;
#pragma unroll
                for (int j = 0; j < VP / 32; ++j) wv[j] = rg_ld_word(lg + 32 * j + lane);
#pragma unroll
                for (int j = 0; j < VP / 32; ++j) {
                    all = all && (32 * j + lane >= p.V || wv[j].y == (unsigned)tick);
                    x[j] = 32 * j + lane < p.V ? __uint_as_float(wv[j].x) : -INFINITY;
                }
                if (__all_sync(0xffffffffu, all ? 1 : 0)) break;
                if (++n > RG_SPIN || ((n & 0x3FFu) == 0 && *(volatile int *)p.err != 0)) { atomicCAS(p.err, 0, 32); alive = false; }
                alive = __all_sync(0xffffffffu, alive ? 1 : 0) != 0;
                if (!alive) break;
            }
            if (!alive) break;
            if (p.debug) { const long long now_ = clock64(); ldbg[0] += now_ - ldbg_t; ldbg_t = now_; }
            float m = -INFINITY, tl = 0.f;
#pragma unroll
            for (int j = 0; j < VP / 32; ++j) {
                if (32 * j + lane == tg) tl = x[j];
                m = fmaxf(m, x[j]);
            }
            for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
            float sum = 0.f;
#pragma unroll
            for (int j = 0; j < VP / 32; ++j) sum += __expf(x[j] - m);                // exp(-inf) = 0 for the padding rows
            for (int o = 16; o > 0; o >>= 1) {
                sum += __shfl_xor_sync(0xffffffffu, sum, o);
                tl += __shfl_xor_sync(0xffffffffu, tl, o);                              // exactly one lane holds the target's logit
            }
            if (p.full_rows) {
                float *row = p.out + (size_t)(o0 + t) * p.V;
                const float inv = 1.0f / sum;
#pragma unroll
                for (int j = 0; j < VP / 32; ++j)
                    if (32 * j + lane < p.V) row[32 * j + lane] = expf(x[j] - m) * inv;
            } else if (lane == 0) {
                p.out[o0 + t] = (tl - m) - logf(sum);
            }
            __syncwarp();
            if (lane == 0) *(volatile int *)lse_done = tick + 1;                        // the slot may be written again
            if (p.debug) { const long long now_ = clock64(); ldbg[1] += now_ - ldbg_t; ldbg_t = now_; }
        }
        if (lane == 0) *(volatile int *)lse_done = 0x7fffffff;
        if (p.debug && lane == 0)
            printf("klm_seq_rg log-sum-exp warp: per tick: waiting for the logits %lld, log-sum-exp %lld clk\n", ldbg[0] / nticks, ldbg[1] / nticks);
    }
    __syncthreads();
    rg_cluster_sync();                            // nobody leaves while a peer may still store into its ring
}

namespace {
template <int WS, int D, int PR, int C> size_t rg_smem_bytes() {
    return (size_t)RG_RING * ((D == 1 ? 1 : (D + 1) / 2 * 2) * 8 * WS + 8 * WS * PR) * 8 + (size_t)2 * D * 8 * WS * 4 + (size_t)(8 * WS / C) * (D + 1) * 16 + 64;
}
template <int WS, int D, int PR, int C> cudaError_t rg_launch(const KlmSeqParams &p, cudaStream_t stream) {
    auto fn = klm_seq_rg<WS, D, PR, C>;
    const size_t smem = rg_smem_bytes<WS, D, PR, C>();
    cudaError_t e = cudaFuncSetAttribute((const void *)fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    if (C > 8) {
        e = cudaFuncSetAttribute((const void *)fn, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        if (e != cudaSuccess) return e;
    }
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.gridDim = dim3(C); cfg.blockDim = dim3(32 * (8 * WS / C + 2)); cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = C; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, fn, p);
}
template <int C> cudaError_t rg_dispatch(const KlmSeqParams &p, cudaStream_t stream) {
    const int PR = p.V <= 128 ? 1 : p.V <= 256 ? 2 : 4;
#define RG_CASE(D_, PR_) if (p.D == D_ && PR == PR_) return rg_launch<16, D_, PR_, C>(p, stream);
    RG_CASE(1, 1) RG_CASE(1, 2) RG_CASE(1, 4) RG_CASE(2, 1) RG_CASE(2, 2) RG_CASE(2, 4) RG_CASE(3, 1) RG_CASE(3, 2) RG_CASE(3, 4)
#undef RG_CASE
    return cudaErrorInvalidValue;
}
}  // namespace

// Models this kernel takes: width 128, depth 1..3, up to 512 characters, ONE stream.
bool klm_rg_eligible(int W, int D, int V) { return W == 128 && D >= 1 && D <= 3 && V >= 2 && V <= 512; }

// cluster: 16 CTAs (KLM_RG_CLUSTER=8 or a refused launch: 8); *cluster_used reports which
cudaError_t klm_launch_seq_rg(const KlmSeqParams &p, cudaStream_t stream, int *cluster) {
    if (*cluster != 8) {
        cudaError_t e = rg_dispatch<16>(p, stream);
        if (e == cudaSuccess) { *cluster = 16; return e; }
        cudaGetLastError();
        *cluster = 8;
    }
    return rg_dispatch<8>(p, stream);
}
