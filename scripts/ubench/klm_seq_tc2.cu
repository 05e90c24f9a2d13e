// CTA-pair variant of the tensor-core sequence kernel (see klm_seq_tc.cu for the algorithm).
//
// The operand fetch of klm_seq_tc is bound by the L2 -> SM port of each SM (~20 B/clk per SM, measured), so
// this kernel halves the bytes every SM needs per tick: the two CTAs of a cluster share one MMA
// (tcgen05.mma.cta_group::2, M = 128 over the pair).  CTA r of the pair holds
//   A rows  : the hidden vector of streams 32r..32r+31 (rows 0..31 fp16 hi part, 32..63 lo part) -- 64 KB at W=512
//   B rows  : its own weight slice (Nc hi columns, Nc lo columns), resident in shared memory as before
// and the tensor cores of both SMs read both B halves.  Two MMAs per k-step: D1 = A . Bhi^T (TMEM columns
// [0, Nc)), D2 = A . Blo^T (columns [Nc, 2Nc)).  In each CTA lanes 0..63 carry the columns of CTA 0's slice and
// lanes 64..127 those of CTA 1's slice (the "2x2" accumulator layout, checked by klm_selftest_umma2), so a CTA
// finishes, for ITS 32 streams, the units of BOTH weight slices of the pair.
//
// Pair protocol: each CTA fetches its own A half (bulk copies, local `full` barriers); CTA 1 relays the
// completion to the leader (`pfull`, remote mbarrier arrive); the leader (CTA 0) issues the MMAs and commits
// with multicast to the `empty` / `tfull` barriers of both CTAs; the epilogue threads of both CTAs arrive on
// the leader's `tempty`.  Everything else (wavefront stages, grid barrier, hand-over buffers) is unchanged.
#include "klm_tc_dev.cuh"

#define TC2_XS (KLM_TC_MAX_PAIRS / 2 * 8 * 32)          // floats of one hand-over direction of one slice

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(KLM_TC_THREADS, 1) klm_seq_tc2(const __grid_constant__ KlmTcParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    const KlmTcRole role = p.roles[blockIdx.x];
    if (role.kind == KLM_TC_LSE) { tc_lse_role(p, role, smem); return; }      // LSE CTAs form clusters of their own

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t rank = tc::cluster_ctarank();
    const int W = p.W, S = p.stages, kc = p.kc;
    const unsigned G = gridDim.x;
    const int nticks = p.T_max + 2 * p.D;
    const int Nc = 8 * role.npm;                            // hi (and lo) columns of each CTA's padded B image
    const int nchunks = W / kc;
    const uint32_t chunk_bytes = (uint32_t)kc * 128;        // kc/8 k-groups x 64 rows x 16 B

    // ---- shared memory carve-up (identical in both CTAs of the pair) ----
    unsigned char *ring = smem;
    float *xchg = reinterpret_cast<float *>(smem);          // aliases the ring (idle while the epilogue runs)
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + p.ring_bytes);   // ring_bytes >= hand-over buffer (20 KB)
    unsigned char *wsm = smem + p.ring_bytes + 256;
    const uint32_t bar_full = tc::smem_u32(bars), bar_empty = tc::smem_u32(bars + 4), bar_pfull = tc::smem_u32(bars + 8);
    const uint32_t bar_tfull = tc::smem_u32(bars + 12), bar_tempty = tc::smem_u32(bars + 13), bar_go = tc::smem_u32(bars + 14);
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 15);

    if (warp == 0) {
        tc::tmem_alloc2(tc::smem_u32(tmem_slot), KLM_TC_TMEM_COLS);
        tc::tmem_relinquish2();
    }
    if (tid == 32) {
        for (int i = 0; i < S; ++i) {
            tc::mbar_init(bar_full + 8 * i, 1); tc::mbar_init(bar_empty + 8 * i, 1); tc::mbar_init(bar_pfull + 8 * i, 1);
        }
        tc::mbar_init(bar_tfull, 1);
        tc::mbar_init(bar_tempty, 512);                      // the epilogue threads of both CTAs
        tc::mbar_init(bar_go, 1);
        tc::fence_mbar_init();
    }
    {   // B image: global -> shared, once per launch
        const uint4 *src = reinterpret_cast<const uint4 *>(p.wimg + role.wofs);
        uint4 *dst = reinterpret_cast<uint4 *>(wsm);
        const int n16 = (W / 8) * 2 * Nc;
        for (int i = tid; i < n16; i += KLM_TC_THREADS) dst[i] = __ldg(src + i);
        tc::fence_proxy_async();
    }

    // ---- epilogue thread identity ----
    // lane quarter q: bit 0 = row half (0: hi rows, 1: lo rows of this CTA's 32 streams), bit 1 = weight slice X
    const int q = warp & 3;
    const int half = q & 1, X = q >> 1;
    const int ch = (warp - 2) >> 2;                          // two warps share a lane quarter and split its pairs
    const int s = 32 * (int)rank + lane;                     // global stream of this thread
    const KlmTcRole roleX = p.roles[blockIdx.x - rank + X];  // the role whose columns this thread finishes
    const int NP = roleX.npairs;
    const int Ph = (NP + 1) / 2;
    const int my_first = half ? Ph : 0, my_n = half ? NP - Ph : Ph;
    const int oth_first = half ? 0 : Ph, oth_n = half ? Ph : NP - Ph;
    const int own_first = my_first + (ch ? (my_n + 1) / 2 : 0), nown = ch ? my_n - (my_n + 1) / 2 : (my_n + 1) / 2;
    const int send_first = oth_first + (ch ? (oth_n + 1) / 2 : 0), nsend = ch ? oth_n - (oth_n + 1) / 2 : (oth_n + 1) / 2;
    const bool is_epi = warp >= 2;
    const bool is_rec = role.kind == KLM_TC_REC;
    const int l = role.layer;
    const size_t hx_half = (size_t)W * 64;                   // halves of one CTA's A image

    float c_reg[KLM_TC_OWN][2], h_reg[KLM_TC_OWN][2];
    long long o0 = 0;
    int len = 0;
    if (is_epi) {
        if (s < p.B) { o0 = p.offsets[s]; len = (int)(p.offsets[s + 1] - o0); }
#pragma unroll
        for (int i = 0; i < KLM_TC_OWN; ++i) {
            c_reg[i][0] = c_reg[i][1] = h_reg[i][0] = h_reg[i][1] = 0.f;
            if (is_rec && i < nown) {
                const int u = 2 * (roleX.first_pair + own_first + i);
                if (s < p.B) {
                    const int slot = p.slots_in[s];
                    if (slot >= 0) {
                        const float *st = p.state + ((size_t)slot * p.D + l) * 2 * W;
                        h_reg[i][0] = st[u]; h_reg[i][1] = st[u + 1];
                        c_reg[i][0] = st[W + u]; c_reg[i][1] = st[W + u + 1];
                    }
                }
                __half hi0, lo0, hi1, lo1;
                tc_split(h_reg[i][0], hi0, lo0); tc_split(h_reg[i][1], hi1, lo1);
                __half *dst = p.hx + ((size_t)(l * 2 + 0) * 2 + rank) * hx_half + (size_t)(u / 8) * (64 * 8) + (u & 7);
                *reinterpret_cast<__half2 *>(dst + lane * 8) = __halves2half2(hi0, hi1);
                *reinterpret_cast<__half2 *>(dst + (32 + lane) * 8) = __halves2half2(lo0, lo1);
            }
        }
        tc::fence_proxy_async();
    }
    tc::tc_fence_before();
    __syncthreads();
    tc::cluster_sync();                                      // the peer's barriers and TMEM are ready, too
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (tid == 0) tc::red_release_add(p.barrier, 1u);

    if (warp == 0) {
        // =============================== producer: this CTA's A half ===============================
        if (lane == 0) {
            uint32_t it = 0;
            bool ok = true;
            for (int tick = 0; tick < nticks && ok; ++tick) {
                if (!tc_wait_grid(p, (unsigned)(tick + 1) * G, 1)) break;
                tc::mbar_arrive(bar_go);
                tc_trace(p, tick, 0);
                tc::fence_proxy_async();                     // generic global writes of other CTAs -> async-proxy reads
                const unsigned char *src = reinterpret_cast<const unsigned char *>(
                    p.hx + ((size_t)(role.src_layer * 2 + (tick & 1)) * 2 + rank) * hx_half);
                for (int c0 = 0; c0 < nchunks; ++c0, ++it) {
                    const int c = (c0 + (int)(blockIdx.x >> 1)) % nchunks;     // same order in both CTAs of the pair
                    const uint32_t st = it % S, ph = (it / S) & 1;
                    if (!tc_wait_mbar(bar_empty + 8 * st, ph ^ 1, p, 2)) { ok = false; break; }
                    tc::mbar_arrive_expect_tx(bar_full + 8 * st, chunk_bytes);
                    tc::bulk_g2s(tc::smem_u32(ring + (size_t)st * chunk_bytes), src + (size_t)c * chunk_bytes, chunk_bytes,
                                 bar_full + 8 * st);
                    if (c0 == 0) tc_trace(p, tick, 1);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && rank == 1) {
            // =============================== relay: tell the leader that our chunks have landed ===============================
            uint32_t it = 0;
            bool ok = true;
            for (int tick = 0; tick < nticks && ok; ++tick)
                for (int c0 = 0; c0 < nchunks; ++c0, ++it) {
                    const uint32_t st = it % S, ph = (it / S) & 1;
                    if (!tc_wait_mbar(bar_full + 8 * st, ph, p, 7)) { ok = false; break; }
                    tc::mbar_arrive_remote(tc::mapa(bar_pfull + 8 * st, 0));
                }
        } else if (lane == 0) {
            // =============================== MMA issuer (leader CTA) ===============================
            const uint32_t idesc = tc::idesc_f16(128, 2 * Nc);
            const uint32_t lbo_a = 64 * 16, lbo_b = 2 * Nc * 16;
            const uint32_t ring_a = tc::smem_u32(ring), w_a = tc::smem_u32(wsm);
            uint32_t it = 0;
            bool ok = true;
            for (int tick = 0; tick < nticks && ok; ++tick) {
                if (!tc_wait_mbar(bar_tempty, (tick & 1) ^ 1, p, 3)) break;
                tc::tc_fence_after();
                for (int c0 = 0; c0 < nchunks; ++c0, ++it) {
                    const int c = (c0 + (int)(blockIdx.x >> 1)) % nchunks;
                    const uint32_t st = it % S, ph = (it / S) & 1;
                    if (!tc_wait_mbar(bar_full + 8 * st, ph, p, 4)) { ok = false; break; }
                    if (!tc_wait_mbar(bar_pfull + 8 * st, ph, p, 8)) { ok = false; break; }
                    tc_trace(p, tick, c0 == 0 ? 2 : 3);
                    tc::tc_fence_after();
                    for (int kk = 0; kk < kc / 16; ++kk) {
                        const uint64_t ad = tc::smem_desc(ring_a + st * chunk_bytes + kk * 2 * lbo_a, lbo_a, 128);
                        const uint32_t boff = w_a + (uint32_t)(c * (kc / 8) + kk * 2) * lbo_b;
                        const uint32_t acc = (c0 | kk) != 0 ? 1u : 0u;
                        tc::mma2_f16(tmem_base, ad, tc::smem_desc(boff, lbo_b, 128), idesc, acc);                  // hi columns
                        tc::mma2_f16(tmem_base + Nc, ad, tc::smem_desc(boff + Nc * 16, lbo_b, 128), idesc, acc);   // lo columns
                    }
                    tc::mma2_commit(bar_empty + 8 * st, 3);        // ring slot free in both CTAs
                }
                if (ok) tc::mma2_commit(bar_tfull, 3);             // accumulators of this tick complete, both CTAs
            }
        }
    } else {
        // =============================== epilogue ===============================
        const uint32_t trow = tmem_base + ((uint32_t)(32 * q) << 16);
        const uint32_t tempty_leader = tc::mapa(bar_tempty, 0);
        float *xs = xchg + (size_t)X * 2 * TC2_XS + (half ? 0 : TC2_XS);        // what this thread hands over
        const float *xr = xchg + (size_t)X * 2 * TC2_XS + (half ? TC2_XS : 0);  // what it receives
        float4 bias4[KLM_TC_OWN][2];
        if (role.kind == KLM_TC_IN) {
#pragma unroll
            for (int i = 0; i < KLM_TC_OWN; ++i)
                if (i < nown) {
                    const int u = 2 * (roleX.first_pair + own_first + i);
                    bias4[i][0] = __ldg(p.bias[l] + u); bias4[i][1] = __ldg(p.bias[l] + u + 1);
                }
        }
        for (int tick = 0; tick < nticks; ++tick) {
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
                    int x = __ldg(p.chars + o0 + t);
                    x = min(max(x, 0), p.V - 1);
                    const bool real = p.n_real ? (t < __ldg(p.n_real + s)) : true;
                    int z[KLM_MAX_CTX];
                    for (int k = 0; k < p.Nc; ++k) {
                        z[k] = real ? __ldg(p.ctx + s * p.Nc + k) : 0;       // padding: context 0 (rating.py:1141-1143)
                        z[k] = min(max(z[k], 0), p.ctx_rows[k] - 1);
                    }
#pragma unroll
                    for (int i = 0; i < KLM_TC_OWN; ++i)
                        if (i < nown) {
                            const int u = 2 * (roleX.first_pair + own_first + i);
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
                            const int u = 2 * (roleX.first_pair + own_first + i);
                            add4[i][0] = __ldcg(zrow + u); add4[i][1] = __ldcg(zrow + u + 1);
                        }
                }
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
#pragma unroll
            for (int i = 0; i < KLM_TC_OWN; ++i)
                if (i < nown) {
                    const int pr = own_first + i;
                    tc::tmem_ld8(trow + 8 * pr, acc[i]);
                    if (half == 0) {
                        float b[8];
                        tc::tmem_ld8(trow + Nc + 8 * pr, b);
                        tc::tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < 8; ++j) acc[i][j] += b[j];
                    } else {
                        tc::tmem_ld_wait();
                    }
                }
#pragma unroll
            for (int i = 0; i < KLM_TC_OWN; ++i)
                if (i < nsend) {
                    const int pr = send_first + i, xi = pr - oth_first;
                    float v[8];
                    tc::tmem_ld8(trow + 8 * pr, v);
                    if (half == 0) {
                        float b[8];
                        tc::tmem_ld8(trow + Nc + 8 * pr, b);
                        tc::tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < 8; ++j) v[j] += b[j];
                    } else {
                        tc::tmem_ld_wait();
                    }
#pragma unroll
                    for (int j = 0; j < 8; ++j) xs[(xi * 8 + j) * 32 + lane] = v[j];
                }
            tc::tc_fence_before();
            if (rank == 0) tc::mbar_arrive(bar_tempty); else tc::mbar_arrive_remote(tempty_leader);
            tc_bar_epi();
#pragma unroll
            for (int i = 0; i < KLM_TC_OWN; ++i)
                if (i < nown) {
                    const int xi = own_first + i - my_first;
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[i][j] += xr[(xi * 8 + j) * 32 + lane];
                }
            // the hand-over buffer aliases the ring: order these generic accesses before the next bulk copies
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");

            // ---- finish and hand over ----
            if (tid == 64) tc_trace(p, tick, 5);
            if (is_rec) {
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
                        const int u = 2 * (roleX.first_pair + own_first + i);
                        __half hi0, lo0, hi1, lo1;
                        tc_split(h_reg[i][0], hi0, lo0); tc_split(h_reg[i][1], hi1, lo1);
                        __half *dst = p.hx + ((size_t)(l * 2 + par_out) * 2 + rank) * hx_half + (size_t)(u / 8) * (64 * 8) + (u & 7);
                        *reinterpret_cast<__half2 *>(dst + lane * 8) = __halves2half2(hi0, hi1);
                        *reinterpret_cast<__half2 *>(dst + (32 + lane) * 8) = __halves2half2(lo0, lo1);
                    }
            } else if (role.kind == KLM_TC_IN) {
                float4 *zrow = reinterpret_cast<float4 *>(p.Z + ((size_t)(l * 2 + par_out) * 64 + s) * 4 * W);
#pragma unroll
                for (int i = 0; i < KLM_TC_OWN; ++i)
                    if (i < nown) {
                        const int u = 2 * (roleX.first_pair + own_first + i);
#pragma unroll
                        for (int e = 0; e < 2; ++e) {
                            const float4 b = bias4[i][e];
                            __stcg(zrow + u + e, make_float4(acc[i][4 * e] + b.x, acc[i][4 * e + 1] + b.y,
                                                             acc[i][4 * e + 2] + b.z, acc[i][4 * e + 3] + b.w));
                        }
                    }
            } else {   // PROJ
                // logits are [parity][Vp/4][64 streams] float4: coalesced over the streams of a warp (read by the LSE stage)
                float4 *lrow = reinterpret_cast<float4 *>(p.logits) + (size_t)par_out * (p.Vp / 4) * 64 + s;
#pragma unroll
                for (int i = 0; i < KLM_TC_OWN; ++i)
                    if (i < nown) {
                        const int g8 = roleX.first_pair + own_first + i;
                        __stcg(lrow + (size_t)(2 * g8) * 64, make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]));
                        __stcg(lrow + (size_t)(2 * g8 + 1) * 64, make_float4(acc[i][4], acc[i][5], acc[i][6], acc[i][7]));
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
                    const int u = 2 * (roleX.first_pair + own_first + i);
                    st[u] = h_reg[i][0]; st[u + 1] = h_reg[i][1];
                    st[W + u] = c_reg[i][0]; st[W + u + 1] = c_reg[i][1];
                }
        }
    }

    tc::tc_fence_before();
    __syncthreads();
    tc::cluster_sync();
    if (warp == 0) {
        __syncwarp();
        tc::tmem_dealloc2(tmem_base, KLM_TC_TMEM_COLS);
    }
}

cudaError_t klm_launch_seq_tc2(const KlmTcParams &p, int grid, size_t smem, cudaStream_t stream) {
    cudaError_t e = cudaFuncSetAttribute((const void *)klm_seq_tc2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(KLM_TC_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative; attr[0].val.cooperative = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, klm_seq_tc2, p);
}
