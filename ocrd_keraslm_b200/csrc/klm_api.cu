// Host side of libkeraslm_b200.so: the C ABI declared in include/keraslm_b200.h.
// Model preparation (weight packing, pre-contracted embedding tables), launch planning,
// the device state pool and the two entry points of the rating path (klm_rate, klm_step).
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <chrono>
#include <cstddef>
#include <cstring>
#include <string>
#include <vector>
#include <queue>

#include "klm_common.cuh"

cudaError_t klm_launch_seq_fma(const KlmSeqParams &p, int grid, size_t smem, cudaStream_t stream);
cudaError_t klm_launch_seq_tc(const KlmTcParams &p, int grid, size_t smem, cudaStream_t stream);
cudaError_t klm_launch_seq_df(const KlmTcParams &p, int grid, size_t smem, cudaStream_t stream);
cudaError_t klm_launch_seq_tw(const KlmTcParams &p, int grid, size_t smem, cudaStream_t stream);
cudaError_t klm_launch_seq_rg(const KlmSeqParams &p, cudaStream_t stream, int *cluster);
bool klm_rg_eligible(int W, int D, int V);
cudaError_t klm_launch_umma2_selftest(const __half *A, const __half *B, float *D, int N, int K, int *err, cudaStream_t stream);
cudaError_t klm_launch_umma_selftest(const __half *A, const __half *B, float *D, int N, int K, int *err, cudaStream_t stream);

// ------------------------------------------------------------------------------------------
static thread_local std::string g_err;

static int fail(int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define CK(call)                                                                              \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess)                                                                \
            return fail(KLM_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_),    \
                        __FILE__, __LINE__);                                                  \
    } while (0)

extern "C" const char *klm_last_error(void) { return g_err.c_str(); }
extern "C" int klm_abi_sizeof(int what) { return what == 0 ? (int)sizeof(klm_config) : what == 1 ? (int)sizeof(klm_info) : -1; }

// ------------------------------------------------------------------------------------------
struct KlmPlan {
    std::vector<KlmRole> roles;
    KlmRole *d_roles = nullptr;
    int grid = 0;
    int wsmem = 0;
    int stage_floats = 0;
    size_t smem = 0;
    int max_streams = 0;
};

// Plan of the tensor-core sequence kernel (klm_seq_tc.cu): one role per CTA + the B images.
struct KlmTcPlan {
    bool ok = false;
    std::vector<KlmTcRole> roles;
    KlmTcRole *d_roles = nullptr;
    unsigned char *d_wimg = nullptr;
    int grid = 0, stages = 0, Vp = 0;
    size_t smem = 0;
    __half *hx = nullptr;
    float *Z = nullptr, *logits = nullptr;
    int *err = nullptr;
    int ingest_bulk = 0, kc = KLM_TC_KC, ring_bytes = 0;     // CTA-pair kernel
    int pub = 0, n_lse = 0, consumer_fence = 1;              // single-CTA kernel
    int stream_b = 0;                                        // single-CTA kernel, streamed-weights form (klm_seq_tcs)
    unsigned long long *trace = nullptr;
    int trace_t0 = 0;
    // dataflow kernel (klm_seq_df.cu): flags / counters that are never reset, its own exchange buffers
    bool df_ok = false;
    size_t df_smem = 0;
    unsigned int *sync = nullptr;
    unsigned epoch = 0, tick_base = 0;
    __half *hx_df = nullptr, *hx0 = nullptr;
    float *Z_df = nullptr, *logits_df = nullptr;
    int n_proj = 0;
    // wide streamed form (klm_seq_tw.cu): 128 streams per launch, its own exchange buffers
    bool wide_ok = false;
    size_t wide_smem = 0;
    __half *hx_w = nullptr;
    float *Z_w = nullptr, *logits_w = nullptr;
    long long wide_launches = 0;
};

// Run-time switches: read from the environment ONCE (klm_create), changed afterwards with klm_set_option -- no getenv on
// the launch path.
struct KlmKnobs {
    int tc_trace_mode = 0;        // KLM_TC_TRACE_MODE
    int tc_debug_skip = 0;        // KLM_TC_DEBUG_SKIP (timing experiments, traced instantiations only)
    int tc_poll = 0;              // KLM_TC_POLL
    int tc_wide = 1;              // KLM_TC_WIDE: 0 = never the 128-stream kernel (streamed-weights plans)
    int tc_skew = 0;              // KLM_TC_SKEW: 0 = by launch length, 2 / 3 = forced
    int tc_zpub = -1;             // KLM_TC_ZPUB: 2 = IN / PROJ stage + bulk-store their slice also with the early signal
    int tc_pub_release = 1;       // KLM_TC_PUB_RELEASE: 1 = a published slice is signalled with a gpu-scope RELEASE.  0 (a relaxed
                                  //   store / add after cp.async.bulk.wait_group 0) is NOT safe: the flag can become visible before the
                                  //   data of the completed bulk store (measured with the dataflow kernel: scripts/exp/df_race.py)
    int tc_df = 1;                // KLM_TC_DF: dataflow kernel (a flag per hand-over, klm_seq_df.cu) where the plan allows it:
                                  //   0 = never, 1 = for klm_step (beam steps: one launch for up to 256 hypotheses, no pipeline fill),
                                  //   2 = for sequences as well (measured slower than the grid-barrier kernel there: DESIGN 3.2d)
    int tc_consumer_fence = 0;    // KLM_TC_CONSUMER_FENCE: fence.proxy.async between the flag polls and the bulk reads
    int tc_exp = 0;               // KLM_TC_EXP: experiment switches of the dataflow kernel (KlmTcParams::exp)
    int tc_fault = 0;             // KLM_TC_FAULT (test hook: one CTA withholds a hand-over signal; the watchdog must fire)
    unsigned tc_spin_limit = 1u << 24;   // KLM_TC_SPIN_LIMIT
    int ll_debug = 0;             // KLM_LL_DEBUG
    int cl_debug = 0;             // KLM_CL_DEBUG
    int step_timing = 0;          // KLM_STEP_TIMING
    int step_server = 1;          // KLM_STEP_SERVER: klm_beam_edge / klm_beam_lattice keep ONE resident instance of klm_seq_df for all their steps
    int step_mapped = 1;          // KLM_STEP_MAPPED: beam steps read / write mapped host memory (no copies around the kernel)
};
static int env_int(const char *name, int dflt) { const char *v = getenv(name); return v ? atoi(v) : dflt; }
static void read_knobs(KlmKnobs &k) {
    k.tc_trace_mode = env_int("KLM_TC_TRACE_MODE", 0);
    k.tc_debug_skip = env_int("KLM_TC_DEBUG_SKIP", 0);
    k.tc_poll = env_int("KLM_TC_POLL", 0);
    k.tc_skew = env_int("KLM_TC_SKEW", 0);
    k.tc_wide = env_int("KLM_TC_WIDE", 1);
    k.tc_zpub = env_int("KLM_TC_ZPUB", -1);
    k.tc_pub_release = env_int("KLM_TC_PUB_RELEASE", 1);
    k.tc_df = env_int("KLM_TC_DF", 1);
    k.tc_consumer_fence = env_int("KLM_TC_CONSUMER_FENCE", 0);
    k.tc_exp = env_int("KLM_TC_EXP", 0);
    k.tc_fault = env_int("KLM_TC_FAULT", 0);
    if (getenv("KLM_TC_SPIN_LIMIT")) k.tc_spin_limit = (unsigned)strtoul(getenv("KLM_TC_SPIN_LIMIT"), nullptr, 10);
    k.ll_debug = env_int("KLM_LL_DEBUG", 0);
    k.cl_debug = env_int("KLM_CL_DEBUG", 0);
    k.step_timing = env_int("KLM_STEP_TIMING", 0);
    k.step_mapped = env_int("KLM_STEP_MAPPED", 1);
    k.step_server = env_int("KLM_STEP_SERVER", 1);
}

#define KLM_RING 4
struct klm_model {
    KlmKnobs knobs;
    int device = 0;
    klm_config cfg{};
    int W = 0, D = 0, V = 0, V4 = 0, Nc = 0;
    int sm_count = 0;
    size_t smem_optin = 0;
    cudaStream_t own_stream = nullptr, stream = nullptr;
    // device-resident model
    float4 *P = nullptr, *Q[KLM_MAX_CTX] = {}, *Wl[KLM_MAX_DEPTH] = {}, *bias[KLM_MAX_DEPTH] = {}, *Ep = nullptr;
    // plans
    KlmPlan plan_k, plan_b;
    KlmTcPlan tc;                      // plan of the tensor-core sequence kernel
    // scratch
    float *hbuf = nullptr, *logits = nullptr;
    unsigned int *barrier = nullptr;
    // barrier-free K-parallel kernel (klm_seq_fma_ll): rings of {value, tick} words, per-CTA progress, watchdog
    uint2 *hbuf_ll = nullptr, *logits_ll = nullptr;
    unsigned int *progress = nullptr;
    int *ll_err = nullptr;
    size_t hbuf_ll_bytes = 0, logits_ll_bytes = 0;
    int ll_launches = 0;
    KlmClGeom cl{};                 // one-cluster kernel (klm_seq_cl): C == 0 = not eligible
    int cl_launches = 0;
    // klm_step with host arrays: page-locked staging (a copy from/to pageable memory blocks in the driver for ~10 us)
    char *h_step = nullptr;
    size_t h_step_bytes = 0;
    // step server (klm_seq_df resident over the steps of one klm_beam_edge / klm_beam_lattice call)
    struct StepServer {
        bool active = false;
        bool wanted = false;           // a native beam search is running: steps of 9..256 hypotheses go to the resident instance
        unsigned seq = 0;
        KlmStepCmd *cmd_host = nullptr, *cmd_host_dev = nullptr, *cmd_dev = nullptr;   // mapped host block (+ its device address), device copy
        unsigned *done_host = nullptr, *done_host_dev = nullptr;                        // mapped completion words, one per LSE CTA
        long long steps = 0, starts = 0;
        double t_relay = 0, t_device = 0, t_start = 0, t_stop = 0;
    } srv;
    char *h_map = nullptr;             // klm_step, beam steps: mapped page-locked memory the kernel reads / writes directly
    void *d_map = nullptr;
    size_t h_map_bytes = 0;
    int *h_err = nullptr;              // page-locked copies of the device-side watchdog words, fetched with the results
    int bp_cap = 0;
    // state pool
    float *state = nullptr;
    int n_slots = 0;
    // per-launch small arrays (pinned ring + device copy)
    char *h_ring[KLM_RING] = {}, *d_ring[KLM_RING] = {};
    cudaEvent_t ring_ev[KLM_RING] = {};
    size_t ring_bytes = 0;
    int ring_pos = 0;
    // staging for KLM_MEM_HOST position arrays
    int *d_chars = nullptr, *d_targets = nullptr;
    float *d_out = nullptr;
    size_t cap_pos = 0, cap_out = 0;
    void *d_topk = nullptr;            // klm_step_topk: [n][k] indices, [n][k] probabilities
    size_t topk_bytes = 0;
    // stats
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool timed = false;
    long long launches = 0, tc_launches = 0, df_launches = 0;
    // KLM_STEP_TIMING=1: host-clock breakdown of klm_step / klm_beam_edge, printed by klm_destroy
    int timing = 0;
    long long t_steps = 0, t_edges = 0;
    double t_step_total = 0, t_step_submit = 0, t_step_wait = 0, t_step_kernel_ms = 0, t_edge_total = 0;
};

// ------------------------------------------------------------------------------------------
// One-time table construction: P = E.K1[:W] + b1 and Q_k = C_k.K1[W+off_k:...], accumulated in
// double so that the tables are correctly rounded float32 values of the exact contraction.
__global__ void klm_prep_table(const float *__restrict__ A /*[rows][kdim]*/, const float *__restrict__ Kmat /*[.][4W]*/,
                               const float *__restrict__ b /*[4W] or null*/, int rows, int kdim, int krow0, int W,
                               float4 *__restrict__ out /*[rows][W]*/) {
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)rows * W) return;
    const int u = (int)(idx % W), r = (int)(idx / W);
    double acc[4] = {0, 0, 0, 0};
    for (int k = 0; k < kdim; ++k) {
        const double a = A[(size_t)r * kdim + k];
        const float *kr = Kmat + (size_t)(krow0 + k) * 4 * W;
#pragma unroll
        for (int g = 0; g < 4; ++g) acc[g] += a * (double)kr[g * W + u];
    }
    if (b) for (int g = 0; g < 4; ++g) acc[g] += (double)b[g * W + u];
    out[idx] = make_float4((float)acc[0], (float)acc[1], (float)acc[2], (float)acc[3]);
}

__global__ void klm_state_zero(float *state, const int *slots, int n, int stride) {
    const int s = blockIdx.x;
    if (s >= n) return;
    float *dst = state + (size_t)slots[s] * stride;
    for (int i = threadIdx.x; i < stride; i += blockDim.x) dst[i] = 0.f;
}
__global__ void klm_state_move(float *state, const int *src, const int *dst, int n, int stride) {
    const int s = blockIdx.x;
    if (s >= n || src[s] == dst[s]) return;
    const float *a = state + (size_t)src[s] * stride;
    float *d = state + (size_t)dst[s] * stride;
    for (int i = threadIdx.x; i < stride; i += blockDim.x) d[i] = a[i];
}
__global__ void klm_state_gather(const float *state, const int *slots, int n, int stride, float *out, int to_pool) {
    const int s = blockIdx.x;
    if (s >= n) return;
    float *pool = const_cast<float *>(state) + (size_t)slots[s] * stride;
    float *lin = out + (size_t)s * stride;
    for (int i = threadIdx.x; i < stride; i += blockDim.x) {
        if (to_pool) pool[i] = lin[i]; else lin[i] = pool[i];
    }
}
__global__ void klm_state_dist(const float *state, const int *a, const int *b, int n, int n_arrays, int W, int stride,
                               float *out) {
    const int s = blockIdx.x / n_arrays, k = blockIdx.x % n_arrays;
    if (s >= n) return;
    const float *x = state + (size_t)a[s] * stride + (size_t)k * W;
    const float *y = state + (size_t)b[s] * stride + (size_t)k * W;
    float acc = 0.f;
    for (int i = threadIdx.x; i < W; i += blockDim.x) { const float d = x[i] - y[i]; acc = fmaf(d, d, acc); }
    __shared__ float red[32];
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int i = 0; i < (int)(blockDim.x + 31) / 32; ++i) t += red[i];
        out[s * n_arrays + k] = sqrtf(t);
    }
}

// ------------------------------------------------------------------------------------------
// Launch plan: which CTA owns which slice of which stage (see klm_seq_fma.cu).
static int make_plan(klm_model *m, bool kpar, KlmPlan *pl) {
    const int W = m->W, D = m->D;
    struct Seg { int kind, layer, items, K, n; };
    std::vector<Seg> segs;
    for (int l = 0; l < D; ++l) segs.push_back({KLM_ROLE_LSTM, l, W, l ? 2 * W : W, 1});
    segs.push_back({KLM_ROLE_PROJ, 0, m->V4 / 4, W, 1});
    const int avail = m->sm_count - 1;                       // one CTA for the LSE stage
    if (avail < (int)segs.size()) return fail(KLM_EINVAL, "device has too few SMs (%d)", m->sm_count);
    pl->stage_floats = kpar ? std::max(8192, KLM_MODEK_MAXB * 2 * W) : 8192;
    const size_t stage_bytes = (size_t)pl->stage_floats * 4;
    if (stage_bytes + 1024 > m->smem_optin) return fail(KLM_EINVAL, "width %d too large for shared staging", W);
    const size_t budget = m->smem_optin - stage_bytes;
    double total_cost = 0; size_t total_bytes = 0; int total_items = 0;
    for (auto &s : segs) { total_cost += (double)s.items * s.K; total_bytes += (size_t)s.items * s.K * 16; total_items += s.items; }
    int n_target = avail;
    if (kpar) {
        const int by_smem = (int)((total_bytes + budget - 1) / budget) + (int)segs.size();
        // latency path: at most one unit per warp in every stage where the SMs allow it (a second unit per warp doubles
        // the time a CTA spends between reading h and publishing its slice)
        int by_warp = 0;
        for (auto &sg : segs) by_warp += (sg.items + KLM_WARPS - 1) / KLM_WARPS;
        n_target = std::min(avail, std::max(std::max(by_smem, by_warp), (int)segs.size()));
    }
    for (;;) {
        for (auto &s : segs) s.n = 1;
        int used = (int)segs.size();
        while (used < n_target) {                            // greedy: relieve the most loaded stage
            int best = -1; double load = -1;
            for (int i = 0; i < (int)segs.size(); ++i) {
                if (segs[i].n >= segs[i].items) continue;
                double ld = std::ceil((double)segs[i].items / segs[i].n) * segs[i].K;
                if (kpar)                                    // rounds of KLM_WARPS units first, then the contraction length
                    ld = std::ceil(std::ceil((double)segs[i].items / segs[i].n) / KLM_WARPS) * 1e9 + segs[i].K;
                if (ld > load) { load = ld; best = i; }
            }
            if (best < 0) break;
            segs[best].n++; used++;
        }
        bool fits = true;
        for (auto &s : segs) {
            const int per = (s.items + s.n - 1) / s.n;
            if ((size_t)per * s.K * 16 > budget) fits = false;
        }
        if (fits) { pl->wsmem = 1; break; }
        if (n_target < avail) { n_target++; continue; }
        pl->wsmem = 0; break;
    }
    pl->roles.clear();
    int per_max = 1; size_t wbytes_max = 0;
    for (auto &s : segs) {
        const int per = (s.items + s.n - 1) / s.n;
        for (int first = 0; first < s.items; first += per) {
            KlmRole r = {s.kind, s.layer, first, std::min(per, s.items - first)};
            pl->roles.push_back(r);
        }
        per_max = std::max(per_max, per);
        wbytes_max = std::max(wbytes_max, (size_t)per * s.K * 16);
    }
    pl->roles.push_back({KLM_ROLE_LSE, 0, 0, 0});
    pl->grid = (int)pl->roles.size();
    pl->smem = stage_bytes + (pl->wsmem ? wbytes_max : 0);
    if (kpar) {
        if (per_max > KLM_WARPS * KLM_MAX_UNITS_K) return fail(KLM_EINVAL, "model too large for the K-parallel plan");
        pl->max_streams = KLM_MODEK_MAXB;
    } else {
        const int nsg = (KLM_WARPS * KLM_MAX_ITEMS) / per_max;
        if (nsg < 1) return fail(KLM_EINVAL, "model too large for the stream-parallel plan");
        pl->max_streams = std::min(256, 32 * nsg);
    }
    CK(cudaMalloc(&pl->d_roles, pl->roles.size() * sizeof(KlmRole)));
    CK(cudaMemcpy(pl->d_roles, pl->roles.data(), pl->roles.size() * sizeof(KlmRole), cudaMemcpyHostToDevice));
    return KLM_OK;
}

// ------------------------------------------------------------------------------------------
// Plan of the single-CTA tensor-core kernel (klm_seq_tc.cu).  Distributes the columns of R_l (REC), K_l (IN,
// l >= 1) and the rows of E (PROJ) over the SMs in pairs of units (8 columns) and builds each CTA's fp16 hi/lo
// B image in the shared-memory operand layout [W/8][2N][8].
//   * REC slices are whole groups of 8 units (4 pairs) so that the new hidden vector leaves the CTA as whole
//     2 KB blocks of the readers' operand layout (role.pub; see the kernel header).  If that does not fit the
//     SM budget the slices fall back to single pairs with scattered publication (pub = 0).
//   * every role gets the largest operand ring that fits beside its B image: (k per chunk, slots) from
//     (64,4), (64,3), (32,4), (32,3), (32,2).
//   * spare SMs go, one refinement at a time, to the stage with the heaviest per-CTA epilogue (REC counts 1.5x:
//     it also runs the LSTM cell and sits on the recurrence).
// Not eligible (ok stays false, the FMA kernels take over) if W is not a multiple of 64, a weight overflows
// fp16, or the images do not fit the device.
static int make_tc1_plan(klm_model *m, const float *E, const float *const *L, KlmTcPlan &pl) {
    const int W = m->W, D = m->D, V = m->V;
    if (getenv("KLM_DISABLE_TC")) return KLM_OK;
    if (W % 64 != 0 || W < 64) return KLM_OK;
    pl.Vp = (V + 7) / 8 * 8;
    static const int geoms[6][2] = {{128, 3}, {64, 4}, {64, 3}, {32, 4}, {32, 3}, {32, 2}};
    const bool kc128 = !(getenv("KLM_TC_KC128") && atoi(getenv("KLM_TC_KC128")) == 0);
    int g_last = 3;                           // the 8 KB chunk geometries only if nothing else fits (each chunk costs the
                                              // producer and the MMA issuer ~400 clk of mbarrier round trips)
    auto ring_bytes_of = [&](int g) {
        const int stages = std::min(geoms[g][1], W / geoms[g][0]);
        return std::max(stages * geoms[g][0] * 256, KLM_TC1_ALIAS_BYTES);
    };
    auto geom_for = [&](int np) {            // first (largest) geometry that fits beside an image of np pairs
        for (int g = 0; g < g_last; ++g) {
            if (W % geoms[g][0] != 0 || (geoms[g][0] == 128 && !kc128)) continue;
            if ((size_t)ring_bytes_of(g) + 128 + (size_t)32 * W * np <= m->smem_optin) return g;
        }
        return -1;
    };
    int pmax = 0;
    const int sm_avail = getenv("KLM_TC_MAX_CTAS") ? std::min(m->sm_count, atoi(getenv("KLM_TC_MAX_CTAS"))) : m->sm_count;
    int pub_mode = getenv("KLM_TC_PUB") ? atoi(getenv("KLM_TC_PUB")) : 2;
    if (pub_mode < 0 || pub_mode > 2) pub_mode = 2;

    struct Ty { int kind, layer, src, sigma, items, gran, n; double weight; const float *mat; };
    std::vector<Ty> tys;
    bool planned = false;
    // distributes the slices of every stage over the SMs: REC items of gran_rec pairs, IN / PROJ items of gran_oth pairs,
    // at most pmax_ pairs per CTA; refines the stage with the largest weighted slice while SMs are left
    auto plan_with = [&](int gran_rec, int gran_oth, int pmax_) -> bool {
        if (gran_rec > pmax_ || gran_oth > pmax_ || (W / 2) % gran_rec != 0 || (W / 2) % gran_oth != 0 || (pl.Vp / 8) % gran_oth != 0)
            return false;
        tys.clear();
        for (int l = 0; l < D; ++l) {
            if (l > 0) tys.push_back({KLM_TC_IN, l, l - 1, 2 * l - 1, W / 2 / gran_oth, gran_oth, 0, 1.0, L[3 * l]});
            tys.push_back({KLM_TC_REC, l, l, 2 * l, W / 2 / gran_rec, gran_rec, 0, 1.5, L[3 * l + 1]});
        }
        tys.push_back({KLM_TC_PROJ, 0, D - 1, 2 * D - 1, pl.Vp / 8 / gran_oth, gran_oth, 0, 1.0, E});
        int used = 0;
        for (auto &t : tys) { const int cap = pmax_ / t.gran; t.n = (t.items + cap - 1) / cap; used += t.n; }
        if (used + 1 > sm_avail) return false;
        const int lse_want = getenv("KLM_TC_NLSE") ? std::max(1, atoi(getenv("KLM_TC_NLSE"))) : 2;
        int n_lse = std::max(1, std::min(lse_want, sm_avail - used));
        used += n_lse;
        auto load = [](const Ty &t, int n) { return (t.items + n - 1) / n * t.gran; };      // pairs of the largest slice
        for (;;) {
            // candidates in order of decreasing weighted load; take the first whose next refinement fits
            std::vector<int> order(tys.size());
            for (size_t i = 0; i < order.size(); ++i) order[i] = (int)i;
            std::sort(order.begin(), order.end(), [&](int a, int b) {
                return load(tys[a], tys[a].n) * tys[a].weight > load(tys[b], tys[b].n) * tys[b].weight; });
            bool done = true;
            for (int i : order) {
                Ty &t = tys[i];
                const int cur = load(t, t.n);
                if (cur <= t.gran) continue;
                int n2 = t.n + 1;
                while (n2 < t.items && load(t, n2) >= cur) ++n2;
                if (load(t, n2) >= cur || used + (n2 - t.n) > sm_avail) continue;
                used += n2 - t.n; t.n = n2; done = false;
                break;
            }
            if (done) break;
        }
        pl.n_lse = n_lse;
        return true;
    };
    // attempts: (aligned REC slices, 16/32 KB chunks), (aligned, any chunk size), (single pairs, ...), (single pairs, any)
    for (int attempt = (pub_mode ? 0 : 2); attempt < 4 && !planned; ++attempt) {
        const int gran = attempt < 2 ? 4 : 1;                // pairs per indivisible REC item
        g_last = (attempt & 1) ? 6 : 3;
        pmax = 0;
        for (int np = 1; np <= KLM_TC1_MAX_PAIRS; ++np) if (geom_for(np) >= 0) pmax = np;
        if (getenv("KLM_TC_PMAX")) pmax = std::min(pmax, atoi(getenv("KLM_TC_PMAX")));
        if (!plan_with(gran, 1, pmax)) continue;
        pl.pub = attempt < 2 ? pub_mode : 0;
        planned = true;
    }
    // streamed-weights form: the B images stay in global memory; every ring slot carries an A chunk of 64 k and the
    // matching part of the image.  Slices are even numbers of pairs (a slice of more than 16 pairs is multiplied in two
    // MMAs of N = 8 * pairs columns, and N must be a multiple of 16); compile-time ring geometries (3 slots) exist for
    // W = 512, 768, 1024, 1536 and 2048.
    auto stream_slot = [&](int np) { return 64 * 256 + 64 * 32 * np; };
    // in front of the ring: the hi<->lo hand-over buffer and, in REC CTAs, the publication staging of the slice
    auto stream_ring_ofs = [&](int np, bool rec) { return KLM_TC1_PUB_OFS + (rec ? (np + 3) / 4 * 2048 : 0); };
    // (1 KB of static shared memory belongs to the LSE role; a 4th slot fits REC slices of 16 pairs but made the
    //  tick 3 % slower: cfg5 3.99 M -> 3.88 M chars/s)
    const size_t stream_max_stages = 3;
    auto stream_stages = [&](int np, bool rec) {
        return (int)std::min<size_t>(stream_max_stages, (m->smem_optin - 1024 - 128 - stream_ring_ofs(np, rec)) / (size_t)stream_slot(np)); };
    const int nch64 = W / 64;
    const bool stream_geom = W % 64 == 0 && (nch64 == 8 || nch64 == 12 || nch64 == 16 || nch64 == 24 || nch64 == 32);
    if (!planned && stream_geom && !getenv("KLM_DISABLE_TCS")) {
        pl.Vp = (V + 15) / 16 * 16;
        pmax = 0;
        for (int np = 2; np <= KLM_TCS_MAX_PAIRS; np += 2) if (stream_stages(np, true) >= 3) pmax = np;
        if (getenv("KLM_TC_PMAX")) pmax = std::min(pmax, atoi(getenv("KLM_TC_PMAX")));
        if (plan_with(4, 2, pmax)) {
            pl.pub = pub_mode ? pub_mode : 2;
            pl.stream_b = 1;
            planned = true;
        } else {
            pl.Vp = (V + 7) / 8 * 8;
        }
    }
    if (!planned) return KLM_OK;
    // fp16 range check
    {
        float mx = 0.f;
        for (auto &t : tys) {
            const size_t cnt = t.kind == KLM_TC_PROJ ? (size_t)V * W : (size_t)W * 4 * W;
            for (size_t i = 0; i < cnt; ++i) mx = std::max(mx, std::fabs(t.mat[i]));
        }
        if (!(mx < 60000.f)) return KLM_OK;
    }
    std::vector<__half> img;
    size_t smem = 0;
    for (auto &t : tys) {
        const int base = t.items / t.n, rem = t.items % t.n;
        int first = 0;
        // power-of-two scale of this matrix (see KLM_TC_HSCALE): the largest that keeps |w| below KLM_TC_WMAX
        float wmax = 0.f;
        {
            const size_t cnt = t.kind == KLM_TC_PROJ ? (size_t)V * W : (size_t)W * 4 * W;
            for (size_t i = 0; i < cnt; ++i) wmax = std::max(wmax, std::fabs(t.mat[i]));
        }
        int wexp = 0;
        if (wmax > 0.f) wexp = std::min(24, std::max(0, (int)std::floor(std::log2(KLM_TC_WMAX / wmax))));
        const float wscale = std::ldexp(1.0f, wexp);
        for (int c = 0; c < t.n; ++c) {
            const int np = (base + (c < rem ? 1 : 0)) * t.gran;
            const int N = 8 * np;                            // image rows: [0,N) hi, [N,2N) lo
            const int g = pl.stream_b ? 0 : geom_for(np);
            KlmTcRole r;
            memset(&r, 0, sizeof r);
            r.kind = t.kind; r.layer = t.layer; r.src_layer = t.src; r.sigma = t.sigma;
            r.first_pair = first; r.npairs = np; r.wofs = (long long)(img.size() * sizeof(__half)); r.npm = np;
            if (pl.stream_b) {
                const bool rec = t.kind == KLM_TC_REC;
                r.kc = 64; r.stages = stream_stages(np, rec); r.ring_ofs = stream_ring_ofs(np, rec);
                r.ring_bytes = r.ring_ofs + r.stages * stream_slot(np);
            } else {
                r.kc = geoms[g][0]; r.stages = std::min(geoms[g][1], W / r.kc); r.ring_bytes = ring_bytes_of(g);
            }
            r.pub = t.kind == KLM_TC_REC ? pl.pub : 0;
            r.scale = 1.0f / (KLM_TC_HSCALE * wscale);
            // IN / PROJ of the resident form: the slice of Z / of the logits (2 KB per pair) staged behind the hand-over
            // buffer on the idle ring and written with one bulk async store (KLM_TC_ZPUB=0: per-thread stores + release)
            if (t.kind != KLM_TC_REC && !pl.stream_b && pl.pub == 2 && r.ring_bytes >= KLM_TC1_PUB_OFS + np * 2048 &&
                !(getenv("KLM_TC_ZPUB") && atoi(getenv("KLM_TC_ZPUB")) == 0))
                r.pub = 2;
            pl.roles.push_back(r);
            smem = std::max(smem, (size_t)r.ring_bytes + 128 + (pl.stream_b ? 0 : (size_t)32 * W * np));
            const size_t o = img.size();
            img.resize(o + (size_t)W * 2 * N, __float2half_rn(0.f));
            for (int j = 0; j < N; ++j) {
                for (int k = 0; k < W; ++k) {
                    float val;
                    if (t.kind == KLM_TC_PROJ) {
                        const int v = 8 * first + j;
                        val = v < V ? t.mat[(size_t)v * W + k] : 0.f;
                    } else {
                        const int u = 2 * first + j / 4, gate = j % 4;
                        val = t.mat[(size_t)k * 4 * W + gate * W + u];
                    }
                    val *= wscale;
                    const __half hi = __float2half_rn(val);
                    const __half lo = __float2half_rn(val - __half2float(hi));
                    img[o + ((size_t)(k / 8) * 2 * N + j) * 8 + (k % 8)] = hi;
                    img[o + ((size_t)(k / 8) * 2 * N + N + j) * 8 + (k % 8)] = lo;
                }
            }
            first += np;
        }
    }
    for (int i = 0; i < pl.n_lse; ++i) {
        KlmTcRole r;
        memset(&r, 0, sizeof r);
        r.kind = KLM_TC_LSE; r.sigma = 2 * D; r.first_pair = i; r.npairs = pl.n_lse;
        pl.roles.push_back(r);
    }
    pl.grid = (int)pl.roles.size();
    pl.smem = smem;
    if (getenv("KLM_TC_PLAN_DEBUG")) {
        static const char *kn[] = {"REC", "IN", "PROJ", "LSE"};
        for (size_t i = 0; i < tys.size(); ++i)
            fprintf(stderr, "klm tc plan: %s l%d: %d CTAs, largest slice %d pairs\n", kn[tys[i].kind], tys[i].layer, tys[i].n,
                    (tys[i].items + tys[i].n - 1) / tys[i].n * tys[i].gran);
        fprintf(stderr, "klm tc plan: LSE %d CTAs; grid %d, smem %zu, pub %d%s\n", pl.n_lse, pl.grid, pl.smem, pl.pub,
                pl.stream_b ? ", streamed weights" : "");
        for (auto &r : pl.roles)
            if (r.kind != KLM_TC_LSE && r.first_pair == 0)
                fprintf(stderr, "klm tc plan:   %s l%d first CTA: %d pairs, ring %d x %d k (%d B)\n", kn[r.kind], r.layer, r.npairs,
                        r.stages, r.kc, r.ring_bytes);
    }
    // the writers order their stores against the async proxy themselves (bulk stores, or fence.proxy.async before
    // the release), so the readers need no proxy fence of their own; KLM_TC_CONSUMER_FENCE=1 adds one (~600 clk per tick)
    pl.consumer_fence = m->knobs.tc_consumer_fence;
    // ---- dataflow kernel: who waits for whom (klm_seq_df.cu) ----
    if (!pl.stream_b && pl.pub == 2) {
        bool ok = true;
        std::vector<std::vector<int>> rec_of(D), in_of(D);
        std::vector<int> proj;
        for (int i = 0; i < (int)pl.roles.size(); ++i) {
            const KlmTcRole &r = pl.roles[i];
            if (r.kind == KLM_TC_REC) rec_of[r.layer].push_back(i);
            else if (r.kind == KLM_TC_IN) in_of[r.layer].push_back(i);
            else if (r.kind == KLM_TC_PROJ) proj.push_back(i);
        }
        static const int df_geoms[7][2] = {{128, 3}, {64, 3}, {64, 4}, {128, 2}, {128, 1}, {64, 2}, {64, 1}};
        size_t df_smem = 0;
        for (int l = 0; l < D && ok; ++l) if (rec_of[l].size() > 64 || rec_of[l].empty() || in_of[l].size() > 64) ok = false;
        if (proj.size() > 64 || proj.empty() || pl.n_lse * (KLM_TC_THREADS / 32) > 64) ok = false;
        for (int i = 0; i < (int)pl.roles.size() && ok; ++i) {
            KlmTcRole &r = pl.roles[i];
            if (r.kind == KLM_TC_LSE) continue;
            bool geom = false;
            for (auto &g : df_geoms) if (r.kc == g[0] && r.stages == g[1]) geom = true;
            const int nch = W / r.kc;
            if (!geom || nch > 16 || r.npairs > KLM_TC1_MAX_PAIRS) { ok = false; break; }
            const std::vector<int> &src = rec_of[r.src_layer];
            r.n_src = (int)src.size();
            for (int c = 0; c < 16; ++c) r.ch_mask[c] = 0;
            for (int c = 0; c < nch; ++c)
                for (int j = 0; j < (int)src.size(); ++j) {
                    const int u0 = 2 * pl.roles[src[j]].first_pair, u1 = u0 + 2 * pl.roles[src[j]].npairs;
                    if (u0 < (c + 1) * r.kc && u1 > c * r.kc) r.ch_mask[c] |= 1ull << j;
                }
            r.idx = 0; r.n_readers = r.n_zprod = r.n_zcons = 0; r.alias_slot = 0; r.guard = 0;
            if (r.kind == KLM_TC_REC) {
                const std::vector<int> &mine = rec_of[r.layer];
                r.idx = (int)(std::find(mine.begin(), mine.end(), i) - mine.begin());
                r.n_readers = r.layer + 1 < D ? (int)in_of[r.layer + 1].size() : (int)proj.size();
                r.n_zprod = r.layer > 0 ? (int)in_of[r.layer].size() : 0;
                // the chunk with this CTA's own slice is fetched last of the first ring round; its slot holds the hand-over
                // buffer and the publication staging while the ring is idle
                const int c_own = 2 * r.first_pair / r.kc, c_end = (2 * (r.first_pair + r.npairs) - 1) / r.kc;
                const int s_eff = std::min(r.stages, nch);
                const size_t alias_bytes = (size_t)2 * ((r.npairs + 1) / 2) * 2048 + (size_t)(r.npairs / 4) * 2048;
                if (c_own == c_end && alias_bytes <= (size_t)r.kc * 256) {
                    r.alias_slot = s_eff - 1;
                    r.sigma = (c_own - r.alias_slot + nch) % nch;           // chunk rotation: first chunk fetched each tick
                } else {
                    r.guard = 1; r.alias_slot = 0; r.sigma = r.idx % nch;
                    if (alias_bytes > (size_t)r.ring_bytes) ok = false;
                }
                // step mode stages the initial hidden values of all hypothesis groups side by side (the ring is idle then)
                if ((size_t)2 * ((r.npairs + 1) / 2) * 2048 + (size_t)KLM_DF_MAXG * (r.npairs / 4) * 2048 > (size_t)r.ring_bytes) ok = false;
            } else if (r.kind == KLM_TC_IN) {
                r.n_zcons = (int)rec_of[r.layer].size();
                r.idx = (int)(std::find(in_of[r.layer].begin(), in_of[r.layer].end(), i) - in_of[r.layer].begin());
            } else {
                r.idx = (int)(std::find(proj.begin(), proj.end(), i) - proj.begin());
            }
            df_smem = std::max(df_smem, (size_t)r.ring_bytes + 256 + (size_t)32 * W * r.npairs);
        }
        if (ok && df_smem <= m->smem_optin && !getenv("KLM_DISABLE_DF")) {
            pl.df_ok = true;
            pl.df_smem = df_smem;
            pl.n_proj = (int)proj.size();
        }
    }
    // ---- wide streamed form (klm_seq_tw.cu): the same roles and images, 128 streams per launch ----
    if (pl.stream_b && pl.pub == 2 && W % 64 == 0) {
        bool ok = true;
        size_t need = 0;
        const size_t avail = m->smem_optin - 1024;                   // (1 KB of static shared memory: the LSE role)
        for (auto &r : pl.roles) {
            if (r.kind == KLM_TC_LSE) continue;
            const size_t pub = r.kind == KLM_TC_REC ? (size_t)(r.npairs / 4) * 4096 : 0;
            const size_t slot = (size_t)2 * 64 * 256 + (size_t)64 * 32 * r.npairs;
            if (r.npairs < 2 || (r.npairs & 1) || r.npairs > 24 || (r.kind == KLM_TC_REC && (r.npairs & 3))) ok = false;
            if (1024 + pub + 2 * slot > avail) ok = false;           // at least two ring slots
            need = std::max(need, std::min(avail, 1024 + pub + 3 * slot));
        }
        if (ok) { pl.wide_ok = true; pl.wide_smem = need; }
    }
    for (size_t i = 0; i < pl.roles.size(); ++i) pl.roles[i].rot = (int)i;
    CK(cudaMalloc(&pl.d_roles, pl.roles.size() * sizeof(KlmTcRole)));
    CK(cudaMemcpy(pl.d_roles, pl.roles.data(), pl.roles.size() * sizeof(KlmTcRole), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&pl.d_wimg, img.size() * sizeof(__half)));
    CK(cudaMemcpy(pl.d_wimg, img.data(), img.size() * sizeof(__half), cudaMemcpyHostToDevice));
    const size_t hx_bytes = (size_t)D * 2 * W * 128 * sizeof(__half);
    const size_t z_bytes = (size_t)D * 4 * 64 * 4 * W * sizeof(float);       // 4 slots (skew 3; skew 2 uses two)
    const size_t lg_bytes = (size_t)4 * 64 * pl.Vp * sizeof(float);
    CK(cudaMalloc(&pl.hx, hx_bytes)); CK(cudaMemset(pl.hx, 0, hx_bytes));
    CK(cudaMalloc(&pl.Z, z_bytes)); CK(cudaMemset(pl.Z, 0, z_bytes));
    CK(cudaMalloc(&pl.logits, lg_bytes)); CK(cudaMemset(pl.logits, 0, lg_bytes));
    CK(cudaMalloc(&pl.err, 256)); CK(cudaMemset(pl.err, 0, 256));
    if (pl.df_ok) {
        const size_t hxv = (size_t)W * 128 * sizeof(__half);                               // one version of one layer
        const size_t zdf = (size_t)D * KLM_DF_NZ * 2 * W * 64 * sizeof(float4);
        const size_t ldf = (size_t)KLM_DF_NL * 2 * (pl.Vp / 4) * 64 * sizeof(float4);
        CK(cudaMalloc(&pl.sync, KLM_DF_WORDS * sizeof(unsigned))); CK(cudaMemset(pl.sync, 0, KLM_DF_WORDS * sizeof(unsigned)));
        CK(cudaMalloc(&pl.hx_df, (size_t)D * KLM_DF_R * hxv)); CK(cudaMemset(pl.hx_df, 0, (size_t)D * KLM_DF_R * hxv));
        CK(cudaMalloc(&pl.hx0, (size_t)D * KLM_DF_MAXG * hxv)); CK(cudaMemset(pl.hx0, 0, (size_t)D * KLM_DF_MAXG * hxv));
        CK(cudaMalloc(&pl.Z_df, zdf)); CK(cudaMemset(pl.Z_df, 0, zdf));
        CK(cudaMalloc(&pl.logits_df, ldf)); CK(cudaMemset(pl.logits_df, 0, ldf));
    }
    if (pl.wide_ok) {
        const size_t hxw = (size_t)D * 2 * 2 * W * 128 * sizeof(__half);
        const size_t zw = (size_t)D * 2 * W * 128 * sizeof(float4);
        const size_t lw = (size_t)2 * (pl.Vp / 4) * 128 * sizeof(float4);
        CK(cudaMalloc(&pl.hx_w, hxw)); CK(cudaMemset(pl.hx_w, 0, hxw));
        CK(cudaMalloc(&pl.Z_w, zw)); CK(cudaMemset(pl.Z_w, 0, zw));
        CK(cudaMalloc(&pl.logits_w, lw)); CK(cudaMemset(pl.logits_w, 0, lw));
    }
    pl.ok = true;                                            // launched cooperatively: all CTAs are co-resident
    return KLM_OK;
}

template <typename T> static int upload(T **dst, const std::vector<T> &src) {
    CK(cudaMalloc(dst, src.size() * sizeof(T)));
    CK(cudaMemcpy(*dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice));
    return KLM_OK;
}

extern "C" int klm_destroy(klm_model *m) {
    if (!m) return KLM_OK;
    cudaSetDevice(m->device);
    if (m->timing && m->t_steps)
        fprintf(stderr, "klm timing: %lld klm_step calls: %.1f us each (submit %.1f, wait for the device %.1f; kernels %.1f us on the "
                "device); %lld klm_beam_edge calls: %.1f us each, of which klm_step %.1f %%\n", m->t_steps,
                1e6 * m->t_step_total / m->t_steps, 1e6 * m->t_step_submit / m->t_steps, 1e6 * m->t_step_wait / m->t_steps,
                1e3 * m->t_step_kernel_ms / m->t_steps, m->t_edges, m->t_edges ? 1e6 * m->t_edge_total / m->t_edges : 0.0,
                m->t_edge_total > 0 ? 100.0 * m->t_step_total / m->t_edge_total : 0.0);
    if (m->timing && m->srv.steps)
        fprintf(stderr, "klm timing: step server: %lld steps in %lld resident instances; per step on the device: doorbell seen -> "
                "device bell %.1f us, -> scores out %.1f us; start %.1f us, stop %.1f us per instance\n", m->srv.steps, m->srv.starts,
                1e6 * m->srv.t_relay / m->srv.steps, 1e6 * m->srv.t_device / m->srv.steps,
                1e6 * m->srv.t_start / std::max(1LL, m->srv.starts), 1e6 * m->srv.t_stop / std::max(1LL, m->srv.starts));
    if (m->stream) cudaStreamSynchronize(m->stream);
    cudaFree(m->P); cudaFree(m->Ep);
    for (int i = 0; i < KLM_MAX_CTX; ++i) cudaFree(m->Q[i]);
    for (int i = 0; i < KLM_MAX_DEPTH; ++i) { cudaFree(m->Wl[i]); cudaFree(m->bias[i]); }
    cudaFree(m->plan_k.d_roles); cudaFree(m->plan_b.d_roles);
    cudaFree(m->tc.d_roles); cudaFree(m->tc.d_wimg); cudaFree(m->tc.hx); cudaFree(m->tc.Z); cudaFree(m->tc.logits);
    cudaFree(m->tc.err); cudaFree(m->tc.trace);
    cudaFree(m->tc.hx_w); cudaFree(m->tc.Z_w); cudaFree(m->tc.logits_w);
    cudaFree(m->tc.sync); cudaFree(m->tc.hx_df); cudaFree(m->tc.hx0); cudaFree(m->tc.Z_df); cudaFree(m->tc.logits_df);
    cudaFree(m->hbuf); cudaFree(m->logits); cudaFree(m->barrier); cudaFree(m->state);
    cudaFree(m->hbuf_ll); cudaFree(m->logits_ll); cudaFree(m->progress); cudaFree(m->ll_err);
    if (m->h_step) cudaFreeHost(m->h_step);
    if (m->h_map) cudaFreeHost(m->h_map);
    if (m->srv.cmd_host) cudaFreeHost(m->srv.cmd_host);
    if (m->srv.done_host) cudaFreeHost(m->srv.done_host);
    cudaFree(m->srv.cmd_dev);
    if (m->h_err) cudaFreeHost(m->h_err);
    for (int i = 0; i < KLM_RING; ++i) {
        if (m->h_ring[i]) cudaFreeHost(m->h_ring[i]);
        cudaFree(m->d_ring[i]);
        if (m->ring_ev[i]) cudaEventDestroy(m->ring_ev[i]);
    }
    cudaFree(m->d_chars); cudaFree(m->d_targets); cudaFree(m->d_out); cudaFree(m->d_topk);
    if (m->ev0) cudaEventDestroy(m->ev0);
    if (m->ev1) cudaEventDestroy(m->ev1);
    if (m->own_stream) cudaStreamDestroy(m->own_stream);
    delete m;
    return KLM_OK;
}

extern "C" int klm_create(const klm_config *cfg, const float *const *weights, int device, klm_model **out) {
    if (!cfg || !weights || !out) return fail(KLM_EINVAL, "klm_create: NULL argument");
    if (cfg->width < 4 || cfg->width % 4 || cfg->depth < 1 || cfg->depth > KLM_MAX_DEPTH || cfg->voc_size < 2 ||
        cfg->n_ctx < 0 || cfg->n_ctx > KLM_MAX_CTX)
        return fail(KLM_EINVAL, "klm_create: unsupported shape width=%d depth=%d voc=%d n_ctx=%d (width must be a "
                    "multiple of 4, depth <= %d, n_ctx <= %d)", cfg->width, cfg->depth, cfg->voc_size, cfg->n_ctx,
                    KLM_MAX_DEPTH, KLM_MAX_CTX);
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail(KLM_EINVAL, "klm_create: no CUDA device %d (have %d)", device, ndev);
    CK(cudaSetDevice(device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) return fail(KLM_EINVAL, "klm_create: device %d is sm_%d%d; this library is built for sm_100a only",
                                     device, prop.major, prop.minor);
    klm_model *m = new klm_model();
    m->device = device; m->cfg = *cfg;
    read_knobs(m->knobs);
    const int W = m->W = cfg->width, D = m->D = cfg->depth, V = m->V = cfg->voc_size, Nc = m->Nc = cfg->n_ctx;
    m->V4 = (V + 3) / 4 * 4;
    m->sm_count = prop.multiProcessorCount;
    m->smem_optin = prop.sharedMemPerBlockOptin;
    int rc = KLM_OK;
#define TRY(x) do { rc = (x); if (rc != KLM_OK) { klm_destroy(m); return rc; } } while (0)
#define TRYC(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { rc = fail(KLM_ECUDA, "%s failed: %s", #call, \
                        cudaGetErrorString(e_)); klm_destroy(m); return rc; } } while (0)
    TRYC(cudaStreamCreateWithFlags(&m->own_stream, cudaStreamNonBlocking));
    m->stream = m->own_stream;
    TRYC(cudaEventCreate(&m->ev0)); TRYC(cudaEventCreate(&m->ev1));

    int ctx_total = 0;
    for (int k = 0; k < Nc; ++k) {
        if (cfg->ctx_rows[k] < 1 || cfg->ctx_dim[k] < 1) { klm_destroy(m); return fail(KLM_EINVAL, "bad context table %d", k); }
        ctx_total += cfg->ctx_dim[k];
    }
    const float *E = weights[0];
    const float *const *C = weights + 1;
    const float *const *L = weights + 1 + Nc;              // K_l, R_l, b_l triples
    const int D1 = W + ctx_total;

    // ---- raw uploads needed by the table kernels ----
    float *dE = nullptr, *dK1 = nullptr, *db1 = nullptr;
    TRYC(cudaMalloc(&dE, (size_t)V * W * 4));  TRYC(cudaMemcpy(dE, E, (size_t)V * W * 4, cudaMemcpyHostToDevice));
    TRYC(cudaMalloc(&dK1, (size_t)D1 * 4 * W * 4)); TRYC(cudaMemcpy(dK1, L[0], (size_t)D1 * 4 * W * 4, cudaMemcpyHostToDevice));
    TRYC(cudaMalloc(&db1, (size_t)4 * W * 4)); TRYC(cudaMemcpy(db1, L[2], (size_t)4 * W * 4, cudaMemcpyHostToDevice));
    TRYC(cudaMalloc(&m->P, (size_t)V * W * 16));
    {
        const long long n = (long long)V * W;
        klm_prep_table<<<(unsigned)((n + 255) / 256), 256, 0, m->stream>>>(dE, dK1, db1, V, W, 0, W, m->P);
        m->launches++;
    }
    int off = W;
    for (int k = 0; k < Nc; ++k) {
        float *dC = nullptr;
        const size_t nb = (size_t)cfg->ctx_rows[k] * cfg->ctx_dim[k] * 4;
        TRYC(cudaMalloc(&dC, nb)); TRYC(cudaMemcpy(dC, C[k], nb, cudaMemcpyHostToDevice));
        TRYC(cudaMalloc(&m->Q[k], (size_t)cfg->ctx_rows[k] * W * 16));
        const long long n = (long long)cfg->ctx_rows[k] * W;
        klm_prep_table<<<(unsigned)((n + 255) / 256), 256, 0, m->stream>>>(dC, dK1, nullptr, cfg->ctx_rows[k],
                                                                            cfg->ctx_dim[k], off, W, m->Q[k]);
        m->launches++;
        TRYC(cudaStreamSynchronize(m->stream));
        cudaFree(dC);
        off += cfg->ctx_dim[k];
    }
    TRYC(cudaStreamSynchronize(m->stream));
    cudaFree(dE); cudaFree(dK1); cudaFree(db1);

    // ---- packed recurrent / input weights: Wl[l][u][k] = (w_i, w_f, w_g, w_o) ----
    for (int l = 0; l < D; ++l) {
        const float *Kl = L[3 * l], *Rl = L[3 * l + 1], *bl = L[3 * l + 2];
        const int Kdim = l ? 2 * W : W;
        std::vector<float4> pk((size_t)W * Kdim);
        for (int u = 0; u < W; ++u)
            for (int k = 0; k < Kdim; ++k) {
                const float *row = (l && k < W) ? Kl + (size_t)k * 4 * W : Rl + (size_t)(l ? k - W : k) * 4 * W;
                pk[(size_t)u * Kdim + k] = make_float4(row[u], row[W + u], row[2 * W + u], row[3 * W + u]);
            }
        TRY(upload(&m->Wl[l], pk));
        std::vector<float4> bb(W);
        for (int u = 0; u < W; ++u) bb[u] = make_float4(bl[u], bl[W + u], bl[2 * W + u], bl[3 * W + u]);
        TRY(upload(&m->bias[l], bb));
    }
    {
        std::vector<float4> pe((size_t)(m->V4 / 4) * W);
        for (int g = 0; g < m->V4 / 4; ++g)
            for (int k = 0; k < W; ++k) {
                float v[4];
                for (int j = 0; j < 4; ++j) v[j] = (4 * g + j < V) ? E[(size_t)(4 * g + j) * W + k] : 0.f;
                pe[(size_t)g * W + k] = make_float4(v[0], v[1], v[2], v[3]);
            }
        TRY(upload(&m->Ep, pe));
    }
    TRY(make_plan(m, true, &m->plan_k));
    TRY(make_plan(m, false, &m->plan_b));
    TRY(make_tc1_plan(m, E, L, m->tc));

    m->bp_cap = std::max(32, m->plan_b.max_streams);
    TRYC(cudaMalloc(&m->hbuf, (size_t)D * 2 * W * m->bp_cap * 4));
    TRYC(cudaMemset(m->hbuf, 0, (size_t)D * 2 * W * m->bp_cap * 4));
    TRYC(cudaMalloc(&m->logits, (size_t)2 * m->V4 * m->bp_cap * 4));
    TRYC(cudaMemset(m->logits, 0, (size_t)2 * m->V4 * m->bp_cap * 4));
    m->timing = m->knobs.step_timing;
    if (!getenv("KLM_DISABLE_LL")) {
        m->hbuf_ll_bytes = (size_t)KLM_LL_RING * D * W * KLM_MODEK_MAXB * sizeof(uint2);
        m->logits_ll_bytes = (size_t)KLM_LL_RING * 4 * m->V4 * KLM_MODEK_MAXB * sizeof(uint2);
        TRYC(cudaMalloc(&m->hbuf_ll, m->hbuf_ll_bytes));
        TRYC(cudaMalloc(&m->logits_ll, m->logits_ll_bytes));
        TRYC(cudaMalloc(&m->progress, 1024 * sizeof(unsigned int)));
        TRYC(cudaMalloc(&m->ll_err, 256));
        TRYC(cudaMemset(m->ll_err, 0, 256));
    }
    // one text at a time, small model: the register-resident cluster kernel (klm_seq_rg.cu); KLM_DISABLE_RG=1 switches it off
    m->cl.C = (m->ll_err && klm_rg_eligible(W, D, V) && !getenv("KLM_DISABLE_RG")) ? (getenv("KLM_RG_CLUSTER") && atoi(getenv("KLM_RG_CLUSTER")) == 8 ? 8 : 16) : 0;
    TRYC(cudaMalloc(&m->barrier, 256));
    m->ring_bytes = (size_t)m->bp_cap * (4 * 4 + 8 + 4 * KLM_MAX_CTX) + 64;
    m->ring_bytes = std::max(m->ring_bytes, (size_t)1 << 16);
    if (m->tc.wide_ok) m->ring_bytes = std::max(m->ring_bytes, (size_t)KLM_TW_MAX_STREAMS * (4 * 4 + 8 + 4 * KLM_MAX_CTX) + 1024);
    for (int i = 0; i < KLM_RING; ++i) {
        TRYC(cudaMallocHost(&m->h_ring[i], m->ring_bytes));
        TRYC(cudaMalloc(&m->d_ring[i], m->ring_bytes));
        TRYC(cudaEventCreateWithFlags(&m->ring_ev[i], cudaEventDisableTiming));
    }
    *out = m;
    return KLM_OK;
}

// ------------------------------------------------------------------------------------------
static int check_device_error(klm_model *m);

extern "C" int klm_set_stream(klm_model *m, void *s) {
    if (!m) return fail(KLM_EINVAL, "NULL model");
    m->stream = s ? (cudaStream_t)s : m->own_stream;
    return KLM_OK;
}
extern "C" int klm_sync(klm_model *m) {
    if (!m) return fail(KLM_EINVAL, "NULL model");
    CK(cudaSetDevice(m->device));
    CK(cudaStreamSynchronize(m->stream));
    return check_device_error(m);
}

extern "C" int klm_set_option(klm_model *m, const char *name, int64_t value) {
    if (!m || !name) return fail(KLM_EINVAL, "klm_set_option: NULL argument");
    KlmKnobs &k = m->knobs;
    const std::string n(name);
    if (n == "tc_trace_mode") k.tc_trace_mode = (int)value;
    else if (n == "tc_debug_skip") k.tc_debug_skip = (int)value;
    else if (n == "tc_poll") k.tc_poll = (int)value;
    else if (n == "tc_skew") k.tc_skew = (int)value;
    else if (n == "tc_wide") k.tc_wide = (int)value;
    else if (n == "tc_zpub") k.tc_zpub = (int)value;
    else if (n == "tc_pub_release") k.tc_pub_release = (int)value;
    else if (n == "tc_fault") k.tc_fault = (int)value;
    else if (n == "tc_exp") k.tc_exp = (int)value;
    else if (n == "tc_df") k.tc_df = (int)value;
    else if (n == "tc_consumer_fence") k.tc_consumer_fence = (int)value;
    else if (n == "tc_spin_limit") k.tc_spin_limit = (unsigned)value;
    else if (n == "ll_debug") k.ll_debug = (int)value;
    else if (n == "step_mapped") k.step_mapped = (int)value;
    else if (n == "step_server") k.step_server = (int)value;
    else if (n == "step_timing") { k.step_timing = (int)value; m->timing = (int)value; }
    else return fail(KLM_EINVAL, "klm_set_option: unknown option '%s'", name);
    return KLM_OK;
}

// Small per-call arrays travel through a ring of pinned buffers: one H2D copy per call.
struct RingSlot { char *h, *d; int idx; size_t used; };
static int ring_acquire(klm_model *m, RingSlot *rs, size_t need) {
    if (need > m->ring_bytes) return fail(KLM_EINVAL, "per-call arrays too large (%zu bytes)", need);
    rs->idx = m->ring_pos; m->ring_pos = (m->ring_pos + 1) % KLM_RING;
    CK(cudaEventSynchronize(m->ring_ev[rs->idx]));
    rs->h = m->h_ring[rs->idx]; rs->d = m->d_ring[rs->idx]; rs->used = 0;
    return KLM_OK;
}
template <typename T> static T *ring_put(RingSlot *rs, const T *src, size_t n, T **dev) {
    rs->used = (rs->used + 15) & ~(size_t)15;
    T *h = reinterpret_cast<T *>(rs->h + rs->used);
    if (src) memcpy(h, src, n * sizeof(T));
    *dev = reinterpret_cast<T *>(rs->d + rs->used);
    rs->used += n * sizeof(T);
    return h;
}
static int ring_commit(klm_model *m, RingSlot *rs) {
    CK(cudaMemcpyAsync(rs->d, rs->h, rs->used, cudaMemcpyHostToDevice, m->stream));
    CK(cudaEventRecord(m->ring_ev[rs->idx], m->stream));
    return KLM_OK;
}

static int check_slots(klm_model *m, const int32_t *slots, int n, bool allow_none, const char *what) {
    if (!slots) return fail(KLM_EINVAL, "%s: NULL slot array", what);
    for (int i = 0; i < n; ++i)
        if (slots[i] >= m->n_slots || slots[i] < (allow_none ? -1 : 0))
            return fail(KLM_ESTATE, "%s: slot %d out of range (pool has %d)", what, slots[i], m->n_slots);
    return KLM_OK;
}

extern "C" int klm_state_pool_resize(klm_model *m, int32_t n_slots) {
    if (!m || n_slots < 0) return fail(KLM_EINVAL, "klm_state_pool_resize: bad argument");
    CK(cudaSetDevice(m->device));
    if (n_slots <= m->n_slots) return KLM_OK;                // never shrinks: handles stay valid
    const size_t stride = (size_t)m->D * 2 * m->W * 4;
    float *ns = nullptr;
    CK(cudaMalloc(&ns, stride * n_slots));
    CK(cudaStreamSynchronize(m->stream));
    CK(cudaMemset(ns, 0, stride * n_slots));
    if (m->state) { CK(cudaMemcpy(ns, m->state, stride * m->n_slots, cudaMemcpyDeviceToDevice)); cudaFree(m->state); }
    m->state = ns; m->n_slots = n_slots;
    return KLM_OK;
}

extern "C" int klm_state_reset(klm_model *m, const int32_t *slots, int32_t n) {
    if (!m) return fail(KLM_EINVAL, "NULL model");
    if (n <= 0) return KLM_OK;
    CK(cudaSetDevice(m->device));
    int rc = check_slots(m, slots, n, false, "klm_state_reset"); if (rc) return rc;
    const int stride = m->D * 2 * m->W;
    for (int i0 = 0; i0 < n; i0 += 4096) {
        const int nb = std::min(4096, n - i0);
        RingSlot rs; rc = ring_acquire(m, &rs, (size_t)nb * 4 + 64); if (rc) return rc;
        int *d; ring_put(&rs, slots + i0, nb, &d);
        rc = ring_commit(m, &rs); if (rc) return rc;
        klm_state_zero<<<nb, 256, 0, m->stream>>>(m->state, d, nb, stride);
        m->launches++;
    }
    CK(cudaGetLastError());
    return KLM_OK;
}

extern "C" int klm_state_copy(klm_model *m, const int32_t *src, const int32_t *dst, int32_t n) {
    if (!m) return fail(KLM_EINVAL, "NULL model");
    if (n <= 0) return KLM_OK;
    CK(cudaSetDevice(m->device));
    int rc = check_slots(m, src, n, false, "klm_state_copy"); if (rc) return rc;
    rc = check_slots(m, dst, n, false, "klm_state_copy"); if (rc) return rc;
    const int stride = m->D * 2 * m->W;
    for (int i0 = 0; i0 < n; i0 += 4096) {
        const int nb = std::min(4096, n - i0);
        RingSlot rs; rc = ring_acquire(m, &rs, (size_t)nb * 8 + 64); if (rc) return rc;
        int *ds, *dd; ring_put(&rs, src + i0, nb, &ds); ring_put(&rs, dst + i0, nb, &dd);
        rc = ring_commit(m, &rs); if (rc) return rc;
        klm_state_move<<<nb, 256, 0, m->stream>>>(m->state, ds, dd, nb, stride);
        m->launches++;
    }
    CK(cudaGetLastError());
    return KLM_OK;
}

static int state_xfer(klm_model *m, const int32_t *slots, int n, float *host, int to_pool) {
    if (!m || !host) return fail(KLM_EINVAL, "state transfer: NULL argument");
    if (n <= 0) return KLM_OK;
    CK(cudaSetDevice(m->device));
    int rc = check_slots(m, slots, n, false, "state transfer"); if (rc) return rc;
    const int stride = m->D * 2 * m->W;
    float *lin = nullptr;
    CK(cudaMalloc(&lin, (size_t)n * stride * 4));
    int *dsl = nullptr;
    CK(cudaMalloc(&dsl, (size_t)n * 4));
    CK(cudaMemcpyAsync(dsl, slots, (size_t)n * 4, cudaMemcpyHostToDevice, m->stream));
    if (to_pool) CK(cudaMemcpyAsync(lin, host, (size_t)n * stride * 4, cudaMemcpyHostToDevice, m->stream));
    klm_state_gather<<<n, 256, 0, m->stream>>>(m->state, dsl, n, stride, lin, to_pool);
    m->launches++;
    if (!to_pool) CK(cudaMemcpyAsync(host, lin, (size_t)n * stride * 4, cudaMemcpyDeviceToHost, m->stream));
    CK(cudaStreamSynchronize(m->stream));
    cudaFree(lin); cudaFree(dsl);
    return KLM_OK;
}
extern "C" int klm_state_to_host(klm_model *m, const int32_t *slots, int32_t n, float *out) {
    return state_xfer(m, slots, n, out, 0);
}
extern "C" int klm_state_from_host(klm_model *m, const int32_t *slots, int32_t n, const float *in) {
    return state_xfer(m, slots, n, const_cast<float *>(in), 1);
}

extern "C" int klm_state_l2_distance(klm_model *m, const int32_t *a, const int32_t *b, int32_t n, int32_t n_arrays,
                                     float *out) {
    if (!m || !out) return fail(KLM_EINVAL, "klm_state_l2_distance: NULL argument");
    if (n <= 0) return KLM_OK;
    if (n_arrays < 1 || n_arrays > 2 * m->D) return fail(KLM_EINVAL, "klm_state_l2_distance: n_arrays %d", n_arrays);
    CK(cudaSetDevice(m->device));
    int rc = check_slots(m, a, n, false, "klm_state_l2_distance"); if (rc) return rc;
    rc = check_slots(m, b, n, false, "klm_state_l2_distance"); if (rc) return rc;
    int *da = nullptr; float *dout = nullptr;
    CK(cudaMalloc(&da, (size_t)n * 8)); CK(cudaMalloc(&dout, (size_t)n * n_arrays * 4));
    CK(cudaMemcpyAsync(da, a, (size_t)n * 4, cudaMemcpyHostToDevice, m->stream));
    CK(cudaMemcpyAsync(da + n, b, (size_t)n * 4, cudaMemcpyHostToDevice, m->stream));
    klm_state_dist<<<n * n_arrays, 128, 0, m->stream>>>(m->state, da, da + n, n, n_arrays, m->W, m->D * 2 * m->W, dout);
    m->launches++;
    CK(cudaMemcpyAsync(out, dout, (size_t)n * n_arrays * 4, cudaMemcpyDeviceToHost, m->stream));
    CK(cudaStreamSynchronize(m->stream));
    cudaFree(da); cudaFree(dout);
    return KLM_OK;
}

// ------------------------------------------------------------------------------------------
// klm_state_l2_distance without any allocation (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: graph capture, tight serving loops).  Same kernel, the scratch
// (device copies of the two slot lists + the distances) in a file-scope device array of KLM_NOALLOC_SCRATCH_BYTES
// that every GPU which loads the library holds until exit.  In host code the array's name is only a stand-in; its
// device address comes from cudaGetSymbolAddress, once per device, with that device current.
// ALL calls share the array (and it keeps the last call's contents): calls that overlap on one GPU race.
static __device__ __align__(256) unsigned char klm_noalloc_scratch[KLM_NOALLOC_SCRATCH_BYTES];
#define KLM_NOALLOC_MAX_DEVICES 64
static void *g_noalloc_ptr[KLM_NOALLOC_MAX_DEVICES];          // device address of klm_noalloc_scratch per device, 0 = not asked yet

static int noalloc_scratch(int device, unsigned char **ptr) {
    if (device < 0 || device >= KLM_NOALLOC_MAX_DEVICES) return fail(KLM_EINVAL, "no-alloc scratch: device %d", device);
    void *p = __atomic_load_n(&g_noalloc_ptr[device], __ATOMIC_ACQUIRE);
    if (!p) {                                                 // set-up, once per device (the device is current: caller)
        const cudaError_t e = cudaGetSymbolAddress(&p, klm_noalloc_scratch);
        if (e != cudaSuccess || !p)
            return fail(KLM_ECUDA, "cudaGetSymbolAddress(klm_noalloc_scratch) on device %d: %s", device, cudaGetErrorString(e));
        __atomic_store_n(&g_noalloc_ptr[device], p, __ATOMIC_RELEASE);   // (two threads racing here store the same address)
    }
    *ptr = static_cast<unsigned char *>(p);
    return KLM_OK;
}

extern "C" int klm_state_l2_distance_noalloc(klm_model *m, const int32_t *a, const int32_t *b, int32_t n, int32_t n_arrays,
                                             float *out) {
    if (!m || !a || !b || !out) return fail(KLM_EINVAL, "klm_state_l2_distance_noalloc: NULL argument");
    if (n <= 0) return KLM_OK;
    if (n_arrays < 1 || n_arrays > 2 * m->D) return fail(KLM_EINVAL, "klm_state_l2_distance_noalloc: n_arrays %d", n_arrays);
    // layout of the scratch: [a: n ints][b: n ints] padded to 256 bytes, then [n][n_arrays] floats
    const size_t slots_bytes = (((size_t)n * 8) + 255) & ~(size_t)255;
    const size_t need = slots_bytes + (size_t)n * n_arrays * 4;
    if (need > (size_t)KLM_NOALLOC_SCRATCH_BYTES)
        return fail(KLM_EINVAL, "klm_state_l2_distance_noalloc: %d pairs x %d arrays need %zu bytes of scratch, the static "
                    "array has %d (use klm_state_l2_distance)", n, n_arrays, need, KLM_NOALLOC_SCRATCH_BYTES);
    CK(cudaSetDevice(m->device));
    int rc = check_slots(m, a, n, false, "klm_state_l2_distance_noalloc"); if (rc) return rc;
    rc = check_slots(m, b, n, false, "klm_state_l2_distance_noalloc"); if (rc) return rc;
    unsigned char *scratch = nullptr;
    rc = noalloc_scratch(m->device, &scratch); if (rc) return rc;
    int *da = reinterpret_cast<int *>(scratch);
    float *dout = reinterpret_cast<float *>(scratch + slots_bytes);
    CK(cudaMemcpyAsync(da, a, (size_t)n * 4, cudaMemcpyHostToDevice, m->stream));
    CK(cudaMemcpyAsync(da + n, b, (size_t)n * 4, cudaMemcpyHostToDevice, m->stream));
    // the kernel writes every distance it is asked for; the clear keeps a short launch from handing out the last call's
    CK(cudaMemsetAsync(dout, 0, (size_t)n * n_arrays * 4, m->stream));
    klm_state_dist<<<n * n_arrays, 128, 0, m->stream>>>(m->state, da, da + n, n, n_arrays, m->W, m->D * 2 * m->W, dout);
    CK(cudaGetLastError());
    m->launches++;
    CK(cudaMemcpyAsync(out, dout, (size_t)n * n_arrays * 4, cudaMemcpyDeviceToHost, m->stream));
    CK(cudaStreamSynchronize(m->stream));
    return KLM_OK;
}

// Fields of the dataflow kernel (klm_seq_df.cu) that do not depend on the batch.
static void tc_df_params(klm_model *m, KlmTcParams &q) {
    KlmTcPlan &tp = m->tc;
    q.sync = tp.sync; q.epoch = tp.epoch; q.tick_base = tp.tick_base;
    q.hx = tp.hx_df; q.hx0 = tp.hx0; q.Z = tp.Z_df; q.logits = tp.logits_df;
    q.n_proj = tp.n_proj; q.n_lse_warps = tp.n_lse * (KLM_TC_THREADS / 32);
    q.consumer_fence = m->knobs.tc_consumer_fence;
    q.exp = m->knobs.tc_exp;
    q.step_groups = 0; q.n_hyp = 0;
}

// After a device-side watchdog error the hand-over words are in an unknown state: start over from zero.
static void tc_df_reset(klm_model *m) {
    KlmTcPlan &tp = m->tc;
    if (tp.sync) cudaMemset(tp.sync, 0, KLM_DF_WORDS * sizeof(unsigned));
    tp.epoch = tp.tick_base = 0;
    if (tp.err) cudaMemset(tp.err, 0, 256);
}

// ------------------------------------------------------------------------------------------
// One cooperative launch over nb streams (nb <= plan.max_streams).  offsets0 are relative to
// d_chars / d_targets / d_out.
static int run_batch(klm_model *m, int nb, const int32_t *slots_in, const int32_t *slots_out, const int64_t *offsets0,
                     const int32_t *n_real, const int32_t *ctx, const int *d_chars, const int *d_targets, float *d_out,
                     int full_rows) {
    KlmPlan &pl = (nb <= KLM_MODEK_MAXB) ? m->plan_k : m->plan_b;
    const bool wide = m->tc.ok && m->tc.wide_ok && nb > KLM_TC_STREAMS && nb <= KLM_TW_MAX_STREAMS;
    if (nb > pl.max_streams && !wide) return fail(KLM_EINVAL, "internal: batch of %d streams exceeds plan (%d)", nb, pl.max_streams);
    long long tmax = 0;
    for (int s = 0; s < nb; ++s) {
        const long long len = offsets0[s + 1] - offsets0[s];
        if (len < 0 || len > 0x7fffffff) return fail(KLM_EINVAL, "bad offsets at stream %d", s);
        tmax = std::max(tmax, len);
    }
    if (tmax == 0) return KLM_OK;
    RingSlot rs;
    int rc = ring_acquire(m, &rs, (size_t)nb * (4 * 4 + 4 * m->Nc) + (nb + 1) * 8 + 130 * 4 + 192);
    if (rc) return rc;
    // wide form: the streams are queued on the 128 rows of the accumulator (longest-processing-time first; a text costs its
    // positions + one gap tick), a row rates its texts one after the other in ONE launch
    std::vector<int> lane_ptr, lane_streams;
    if (wide) {
        const int R = 2 * KLM_TC_STREAMS;
        std::vector<int> order(nb);
        for (int i = 0; i < nb; ++i) order[i] = i;
        std::stable_sort(order.begin(), order.end(), [&](int a, int b) {
            return offsets0[a + 1] - offsets0[a] > offsets0[b + 1] - offsets0[b]; });
        typedef std::pair<long long, int> LR;                                   // (load, row): least loaded first, ties by row
        std::priority_queue<LR, std::vector<LR>, std::greater<LR>> heap;
        for (int r = 0; r < R; ++r) heap.push(LR(0, r));
        std::vector<std::vector<int>> rows(R);
        for (int i : order) {
            LR top = heap.top(); heap.pop();
            rows[top.second].push_back(i);
            heap.push(LR(top.first + (offsets0[i + 1] - offsets0[i]) + 1, top.second));
        }
        lane_ptr.assign(R + 1, 0);
        tmax = 0;
        for (int r = 0; r < R; ++r) {
            long long load = 0;
            for (int i : rows[r]) { lane_streams.push_back(i); load += (offsets0[i + 1] - offsets0[i]) + 1; }
            lane_ptr[r + 1] = (int)lane_streams.size();
            tmax = std::max(tmax, load);
        }
    }
    if (m->tc.ok && nb > KLM_MODEK_MAXB && (nb <= KLM_TC_STREAMS || wide)) {
        KlmTcPlan &tp = m->tc;
        if ((double)(tmax + 3 * m->D + 2) * tp.grid >= 4.0e9)
            return fail(KLM_EINVAL, "stream of %lld positions is too long for one launch", tmax);
        KlmTcParams q;
        memset(&q, 0, sizeof q);
        int *d_si, *d_so, *d_nr = nullptr, *d_cx;
        long long *d_off;
        ring_put(&rs, slots_in, nb, &d_si);
        ring_put(&rs, slots_out, nb, &d_so);
        ring_put(&rs, reinterpret_cast<const long long *>(offsets0), nb + 1, &d_off);
        if (n_real) ring_put(&rs, n_real, nb, &d_nr);
        int *hcx = ring_put<int>(&rs, nullptr, (size_t)nb * std::max(1, m->Nc), &d_cx);
        for (int i = 0; i < nb * m->Nc; ++i) hcx[i] = ctx ? ctx[i] : 0;
        int *d_lp = nullptr, *d_ls = nullptr;
        if (wide) {
            ring_put(&rs, lane_ptr.data(), lane_ptr.size(), &d_lp);
            ring_put(&rs, lane_streams.data(), lane_streams.size(), &d_ls);
        }
        rc = ring_commit(m, &rs); if (rc) return rc;
        q.lane_ptr = d_lp; q.lane_streams = d_ls;
        q.W = m->W; q.D = m->D; q.V = m->V; q.Vp = tp.Vp; q.Nc = m->Nc;
        q.B = nb; q.T_max = (int)tmax; q.stages = tp.stages;
        for (int k = 0; k < m->Nc; ++k) { q.ctx_rows[k] = m->cfg.ctx_rows[k]; q.Q[k] = m->Q[k]; }
        q.P = m->P;
        for (int l = 0; l < m->D; ++l) q.bias[l] = m->bias[l];
        q.wimg = tp.d_wimg; q.roles = tp.d_roles;
        q.hx = tp.hx; q.Z = tp.Z; q.logits = tp.logits;
        q.ingest_bulk = tp.ingest_bulk; q.consumer_fence = tp.consumer_fence;
        q.state = m->state; q.slots_in = d_si; q.slots_out = d_so; q.offsets = d_off;
        q.chars = d_chars; q.targets = d_targets; q.n_real = d_nr; q.ctx = d_cx;
        q.out = d_out; q.full_rows = full_rows;
        q.barrier = m->barrier; q.err = tp.err;
        q.trace = tp.trace; q.trace_t0 = tp.trace_t0;
        q.trace_mode = m->knobs.tc_trace_mode;
        q.debug_skip = m->knobs.tc_debug_skip;
        q.poll_mode = m->knobs.tc_poll;
        q.stream_b = tp.stream_b;
        // long launches: skew 3 (the recurrence does not wait for the IN / PROJ epilogues; D more ticks of pipeline fill);
        // short ones (beam steps) and the CTA-pair kernel: skew 2
        {
            // (streamed-weights form: measured slower with skew 3 on the short lines of cfg5, 3.73 M against 3.90 M chars/s)
            const int want = m->knobs.tc_skew ? m->knobs.tc_skew : (tmax >= 64 && !tp.stream_b ? 3 : 2);
            q.skew = want == 3 ? 3 : 2;
            q.zmask = q.skew == 3 ? 3 : 1; q.zlag = q.skew == 3 ? 2 : 1;
        }
        q.zpub_early = m->knobs.tc_zpub == 2;
        q.pub_release = m->knobs.tc_pub_release;
        q.spin_limit = m->knobs.tc_spin_limit;
        q.fault = m->knobs.tc_fault;
        q.kc = tp.kc; q.ring_bytes = tp.ring_bytes;
        if (tp.df_ok && m->knobs.tc_df >= 2) {
            // dataflow kernel: flags / counters continue above the values the earlier launches left (no memset)
            tc_df_params(m, q);
            q.T_max = (int)tmax;
            CK(klm_launch_seq_df(q, tp.grid, tp.df_smem, m->stream));
            tp.tick_base += (unsigned)tmax; tp.epoch += (unsigned)tmax + 1u;
            m->launches++; m->tc_launches++; m->df_launches++;
            return KLM_OK;
        }
        CK(cudaMemsetAsync(m->barrier, 0, 4, m->stream));
        if (wide) {
            // 65..128 streams of a streamed-weights plan: one launch of klm_seq_tw (skew 2, two slots of Z / logits)
            q.wide = 1; q.hx = tp.hx_w; q.Z = tp.Z_w; q.logits = tp.logits_w; q.exp = m->knobs.tc_exp;
            q.B = std::min(nb, 2 * KLM_TC_STREAMS);                          // rows in use
            q.skew = 2; q.zmask = 1; q.zlag = 1;
            q.ring_bytes = (int)tp.wide_smem;
            CK(klm_launch_seq_tw(q, tp.grid, tp.wide_smem, m->stream));
            m->launches++; m->tc_launches++; tp.wide_launches++;
            return KLM_OK;
        }
        CK(klm_launch_seq_tc(q, tp.grid, tp.smem, m->stream));
        m->launches++; m->tc_launches++;
        return KLM_OK;
    }
    KlmSeqParams p;
    memset(&p, 0, sizeof p);
    int *d_si, *d_so, *d_nr = nullptr, *d_cx;
    long long *d_off;
    ring_put(&rs, slots_in, nb, &d_si);
    ring_put(&rs, slots_out, nb, &d_so);
    ring_put(&rs, reinterpret_cast<const long long *>(offsets0), nb + 1, &d_off);
    if (n_real) ring_put(&rs, n_real, nb, &d_nr);
    int *hcx = ring_put<int>(&rs, nullptr, (size_t)nb * std::max(1, m->Nc), &d_cx);
    for (int i = 0; i < nb * m->Nc; ++i) hcx[i] = ctx ? ctx[i] : 0;
    rc = ring_commit(m, &rs); if (rc) return rc;

    p.W = m->W; p.D = m->D; p.V = m->V; p.V4 = m->V4; p.Nc = m->Nc;
    p.B = nb; p.Bp = (nb <= KLM_MODEK_MAXB) ? nb : (nb + 31) / 32 * 32; p.T_max = (int)tmax;
    for (int k = 0; k < m->Nc; ++k) { p.ctx_rows[k] = m->cfg.ctx_rows[k]; p.Q[k] = m->Q[k]; }
    p.P = m->P; p.Ep = m->Ep;
    for (int l = 0; l < m->D; ++l) { p.Wl[l] = m->Wl[l]; p.bias[l] = m->bias[l]; }
    p.state = m->state; p.slots_in = d_si; p.slots_out = d_so; p.offsets = d_off;
    p.chars = d_chars; p.targets = d_targets; p.n_real = d_nr; p.ctx = d_cx;
    p.out = d_out; p.full_rows = full_rows;
    p.hbuf = m->hbuf; p.logits = m->logits; p.barrier = m->barrier;
    if (nb == 1 && m->cl.C > 0 && m->ll_err) {
        // ONE stream, small model: weights in the registers of one thread-block cluster, recurrence through distributed shared memory
        p.err = m->ll_err;
        p.debug = m->knobs.ll_debug;
        cudaError_t e = klm_launch_seq_rg(p, m->stream, &m->cl.C);
        if (e == cudaSuccess) { m->launches++; m->ll_launches++; m->cl_launches++; return KLM_OK; }
        fprintf(stderr, "klm: cluster kernel cannot be launched (%s), using the grid-wide kernel\n", cudaGetErrorString(e));
        cudaGetLastError();
        m->cl.C = 0;
    }
    if (nb <= KLM_MODEK_MAXB && m->hbuf_ll && pl.grid <= 1024) {
        // exchange rings: no word of an earlier launch may look like one of this launch
        p.hbuf_ll = m->hbuf_ll; p.logits_ll = m->logits_ll; p.progress = m->progress; p.err = m->ll_err;
        const size_t used_slots = (size_t)std::min<long long>(KLM_LL_RING, tmax + m->D + 2);   // short launches touch few slots
        CK(cudaMemsetAsync(m->hbuf_ll, 0xFF, m->hbuf_ll_bytes / KLM_LL_RING * used_slots, m->stream));
        CK(cudaMemsetAsync(m->logits_ll, 0xFF, m->logits_ll_bytes / KLM_LL_RING * used_slots, m->stream));
        CK(cudaMemsetAsync(m->progress, 0, 1024 * sizeof(unsigned int), m->stream));
        m->ll_launches++;
        p.debug = m->knobs.ll_debug;
    }
    p.roles = pl.d_roles; p.wsmem = pl.wsmem; p.stage_floats = pl.stage_floats;
    CK(cudaMemsetAsync(m->barrier, 0, 4, m->stream));
    CK(klm_launch_seq_fma(p, pl.grid, pl.smem, m->stream));
    m->launches++;
    return KLM_OK;
}

// Step mode of the dataflow kernel: ONE position for nb <= 64 * KLM_DF_MAXG hypotheses in one launch (tick g advances
// hypotheses 64g .. 64g+63; states in / out through the pool).
struct StepHostArgs { const int32_t *chars, *targets, *slots_in, *slots_out, *ctx; };
static int launch_step_df(klm_model *m, int nb, const int *d_si, const int *d_so, const int *d_cx, const int *d_chars,
                          const int *d_targets, float *d_out, int full_rows, int *err_host, const StepHostArgs *inl = nullptr);

static int run_step_df(klm_model *m, int nb, const int32_t *slots_in, const int32_t *slots_out, const int32_t *ctx,
                       const int *d_chars, const int *d_targets, float *d_out, int full_rows) {
    RingSlot rs;
    int rc = ring_acquire(m, &rs, (size_t)nb * (4 * 2 + 4 * m->Nc) + 128);
    if (rc) return rc;
    int *d_si, *d_so, *d_cx;
    ring_put(&rs, slots_in, nb, &d_si);
    ring_put(&rs, slots_out, nb, &d_so);
    int *hcx = ring_put<int>(&rs, nullptr, (size_t)nb * std::max(1, m->Nc), &d_cx);
    for (int i = 0; i < nb * m->Nc; ++i) hcx[i] = ctx ? ctx[i] : 0;
    rc = ring_commit(m, &rs); if (rc) return rc;
    return launch_step_df(m, nb, d_si, d_so, d_cx, d_chars, d_targets, d_out, full_rows, nullptr);
}

// all pointers: visible to the device (device memory, or mapped page-locked host memory)
static int launch_step_df(klm_model *m, int nb, const int *d_si, const int *d_so, const int *d_cx, const int *d_chars,
                          const int *d_targets, float *d_out, int full_rows, int *err_host, const StepHostArgs *inl) {
    KlmTcPlan &tp = m->tc;
    static thread_local KlmTcParams q;                               // (12 KB: not on the stack of every call)
    memset(&q, 0, offsetof(KlmTcParams, step));
    if (inl) {
        q.step_inline = 1;
        memcpy(q.step.chars, inl->chars, (size_t)nb * 4);
        if (inl->targets) memcpy(q.step.targets, inl->targets, (size_t)nb * 4);
        memcpy(q.step.slots_in, inl->slots_in, (size_t)nb * 4);
        memcpy(q.step.slots_out, inl->slots_out, (size_t)nb * 4);
        for (int i = 0; i < nb * m->Nc; ++i) q.step.ctx[i] = inl->ctx ? inl->ctx[i] : 0;
    }
    q.W = m->W; q.D = m->D; q.V = m->V; q.Vp = tp.Vp; q.Nc = m->Nc;
    q.B = std::min(nb, KLM_TC_STREAMS); q.T_max = 1; q.stages = tp.stages;
    for (int k = 0; k < m->Nc; ++k) { q.ctx_rows[k] = m->cfg.ctx_rows[k]; q.Q[k] = m->Q[k]; }
    q.P = m->P;
    for (int l = 0; l < m->D; ++l) q.bias[l] = m->bias[l];
    q.wimg = tp.d_wimg; q.roles = tp.d_roles;
    q.state = m->state; q.slots_in = d_si; q.slots_out = d_so; q.offsets = nullptr;
    q.chars = d_chars; q.targets = d_targets; q.n_real = nullptr; q.ctx = d_cx;
    q.out = d_out; q.full_rows = full_rows;
    q.barrier = m->barrier; q.err = tp.err;
    q.trace = tp.trace; q.trace_t0 = tp.trace_t0; q.trace_mode = m->knobs.tc_trace_mode;
    q.pub_release = m->knobs.tc_pub_release; q.spin_limit = m->knobs.tc_spin_limit; q.fault = m->knobs.tc_fault;
    tc_df_params(m, q);
    q.step_groups = (nb + KLM_TC_STREAMS - 1) / KLM_TC_STREAMS;
    q.n_hyp = nb;
    q.err_host = err_host;
    CK(klm_launch_seq_df(q, tp.grid, tp.df_smem, m->stream));
    tp.tick_base += (unsigned)q.step_groups; tp.epoch += (unsigned)q.step_groups + 1u;
    m->launches++; m->tc_launches++; m->df_launches++;
    return KLM_OK;
}

// Largest number of streams one launch takes for a call of n streams.
static int batch_max(klm_model *m, int n) {
    if (n <= KLM_MODEK_MAXB) return KLM_MODEK_MAXB;
    if (m->tc.ok) return (m->tc.wide_ok && m->knobs.tc_wide && n > KLM_TC_STREAMS) ? KLM_TW_MAX_STREAMS : KLM_TC_STREAMS;
    return m->plan_b.max_streams;
}

// Device-side watchdog of the tensor-core kernel: a wait that timed out leaves a code in tc.err.
// Error words copied asynchronously in front of a stream synchronisation the caller does anyway.
static int fetch_errors_async(klm_model *m) {
    if (!m->h_err) { CK(cudaHostAlloc(&m->h_err, 64, cudaHostAllocDefault)); }
    m->h_err[0] = m->h_err[1] = m->h_err[2] = 0;
    if (m->ll_launches > 0) CK(cudaMemcpyAsync(&m->h_err[0], m->ll_err, 4, cudaMemcpyDeviceToHost, m->stream));
    if (m->tc.ok && m->tc_launches > 0) {
        CK(cudaMemcpyAsync(&m->h_err[1], m->tc.err, 4, cudaMemcpyDeviceToHost, m->stream));
    }
    return KLM_OK;
}
static int check_fetched_errors(klm_model *m) {
    if (m->h_err[0]) return fail(KLM_ECUDA, "klm_seq_fma_ll: device-side wait timed out (code %d)", m->h_err[0]);
    if (m->h_err[1] || m->h_err[2]) {
        const int code = m->h_err[1] ? m->h_err[1] : m->h_err[2];
        cudaStreamSynchronize(m->stream);
        tc_df_reset(m);                                      // the model stays usable: hand-over words and error word start over
        return fail(KLM_ECUDA, "klm_seq_tc: device-side wait timed out (code %d)", code);
    }
    return KLM_OK;
}

static int check_device_error(klm_model *m) {
    if (m->ll_launches > 0) {
        int code = 0;
        CK(cudaMemcpy(&code, m->ll_err, sizeof code, cudaMemcpyDeviceToHost));
        if (code != 0) return fail(KLM_ECUDA, "klm_seq_fma_ll: device-side wait timed out (code %d)", code);
    }
    if (!m->tc.ok || m->tc_launches == 0) return KLM_OK;
    int code = 0;
    CK(cudaMemcpy(&code, m->tc.err, sizeof code, cudaMemcpyDeviceToHost));
    if (code != 0) {
        cudaStreamSynchronize(m->stream);
        tc_df_reset(m);
        return fail(KLM_ECUDA, "klm_seq_tc: device-side wait timed out (code %d)", code);
    }
    return KLM_OK;
}

static int ensure_pos_cap(klm_model *m, size_t npos, size_t nout) {
    if (npos > m->cap_pos) {
        CK(cudaStreamSynchronize(m->stream));
        cudaFree(m->d_chars); cudaFree(m->d_targets); m->d_chars = m->d_targets = nullptr; m->cap_pos = 0;
        const size_t cap = npos + npos / 4 + 1024;
        CK(cudaMalloc(&m->d_chars, cap * 4)); CK(cudaMalloc(&m->d_targets, cap * 4));
        m->cap_pos = cap;
    }
    if (nout > m->cap_out) {
        CK(cudaStreamSynchronize(m->stream));
        cudaFree(m->d_out); m->d_out = nullptr; m->cap_out = 0;
        const size_t cap = nout + nout / 4 + 1024;
        CK(cudaMalloc(&m->d_out, cap * 4));
        m->cap_out = cap;
    }
    return KLM_OK;
}

extern "C" int klm_rate(klm_model *m, int32_t n, const int32_t *slots, const int64_t *offsets, const int32_t *chars,
                        const int32_t *targets, const int32_t *n_real, const int32_t *ctx, float *out_logp, int mem) {
    if (!m) return fail(KLM_EINVAL, "NULL model");
    if (n <= 0) return KLM_OK;
    if (!offsets || !chars || !targets || !out_logp) return fail(KLM_EINVAL, "klm_rate: NULL argument");
    if (m->Nc > 0 && !ctx) return fail(KLM_EINVAL, "klm_rate: model has %d context variables but ctx is NULL", m->Nc);
    CK(cudaSetDevice(m->device));
    int rc = check_slots(m, slots, n, false, "klm_rate"); if (rc) return rc;
    const long long total = offsets[n] - offsets[0];
    if (total < 0) return fail(KLM_EINVAL, "klm_rate: offsets not increasing");
    if (total == 0) return KLM_OK;
    for (int k = 0; k < m->Nc; ++k)
        for (int s = 0; s < n; ++s)
            if (ctx[s * m->Nc + k] < 0 || ctx[s * m->Nc + k] >= m->cfg.ctx_rows[k])
                return fail(KLM_EINVAL, "klm_rate: context %d of stream %d is %d, table has %d rows", k, s,
                            ctx[s * m->Nc + k], m->cfg.ctx_rows[k]);
    const int *dc, *dt; float *dout;
    const long long base = offsets[0];
    if (mem == KLM_MEM_HOST) {
        rc = ensure_pos_cap(m, (size_t)total, (size_t)total); if (rc) return rc;
        CK(cudaMemcpyAsync(m->d_chars, chars + base, (size_t)total * 4, cudaMemcpyHostToDevice, m->stream));
        CK(cudaMemcpyAsync(m->d_targets, targets + base, (size_t)total * 4, cudaMemcpyHostToDevice, m->stream));
        dc = m->d_chars; dt = m->d_targets; dout = m->d_out;
    } else {
        dc = chars + base; dt = targets + base; dout = out_logp + base;
    }
    CK(cudaEventRecord(m->ev0, m->stream));
    const int bmax = batch_max(m, n);
    std::vector<int64_t> off0;
    for (int s0 = 0; s0 < n; s0 += bmax) {
        const int nb = std::min(bmax, n - s0);
        off0.resize(nb + 1);
        for (int i = 0; i <= nb; ++i) off0[i] = offsets[s0 + i] - base;
        rc = run_batch(m, nb, slots + s0, slots + s0, off0.data(), n_real ? n_real + s0 : nullptr,
                       ctx ? ctx + (size_t)s0 * m->Nc : nullptr, dc, dt, dout, 0);
        if (rc) return rc;
    }
    CK(cudaEventRecord(m->ev1, m->stream));
    m->timed = true;
    if (mem == KLM_MEM_HOST) {
        CK(cudaMemcpyAsync(out_logp + base, m->d_out, (size_t)total * 4, cudaMemcpyDeviceToHost, m->stream));
        CK(cudaStreamSynchronize(m->stream));
        return check_device_error(m);
    }
    return KLM_OK;
}

// k best entries of every softmax row, ascending by (probability, index) -- the order of np.argsort(p)[-k:] for distinct
// values (rating.py:690-692); one warp per row, k rounds of a warp-wide arg-max over what is left.
__global__ void klm_topk_rows(const float *__restrict__ rows, int n, int V, int k, int *__restrict__ out_idx,
                              float *__restrict__ out_prob) {
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= n) return;
    const float *p = rows + (size_t)row * V;
    float last_p = INFINITY;
    int last_i = 0x7fffffff;
    for (int j = 0; j < k; ++j) {
        float best = -1.f;
        int bi = -1;
        for (int v = lane; v < V; v += 32) {
            const float x = p[v];
            // still available = strictly below the last pick in (probability, index) order
            const bool avail = x < last_p || (x == last_p && v < last_i);
            if (avail && (x > best || (x == best && v > bi))) { best = x; bi = v; }
        }
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ob > best || (ob == best && oi > bi)) { best = ob; bi = oi; }
        }
        last_p = best; last_i = bi;
        if (lane == 0) { out_idx[(size_t)row * k + (k - 1 - j)] = bi; out_prob[(size_t)row * k + (k - 1 - j)] = bi >= 0 ? best : 0.f; }
    }
}

// ------------------------------------------------------------------------------------------
// Step server: one resident instance of klm_seq_df serves all beam steps of a klm_beam_edge / klm_beam_lattice call.  A step
// is a KlmStepCmd written into mapped host memory (sequence number last); the scores come back through mapped memory as
// before, completion is a word per LSE CTA.  Nothing else may use the model's stream while the instance is resident (the
// model is not re-entrant, and the two callers do nothing on the device but steps).
static int server_ensure_map(klm_model *m, size_t need) {
    if (need <= m->h_map_bytes) return KLM_OK;
    CK(cudaStreamSynchronize(m->stream));
    if (m->h_map) cudaFreeHost(m->h_map);
    m->h_map = nullptr; m->h_map_bytes = 0;
    CK(cudaHostAlloc(&m->h_map, need * 2, cudaHostAllocMapped));
    CK(cudaHostGetDevicePointer(&m->d_map, m->h_map, 0));
    m->h_map_bytes = need * 2;
    return KLM_OK;
}

static int server_stop(klm_model *m) {
    klm_model::StepServer &sv = m->srv;
    if (!sv.active) return KLM_OK;
    sv.active = false;
    const auto t0 = std::chrono::steady_clock::now();
    sv.cmd_host->n_hyp = -1; sv.cmd_host->groups = 0;
    __atomic_store_n(&sv.cmd_host->seq, ++sv.seq, __ATOMIC_RELEASE);
    CK(cudaStreamSynchronize(m->stream));
    sv.t_stop += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    const int32_t code = *reinterpret_cast<volatile int32_t *>(m->h_map);
    if (code) {
        tc_df_reset(m);
        return fail(KLM_ECUDA, "klm_seq_df (step server): device-side wait timed out (code %d)", code);
    }
    return KLM_OK;
}

// (steps with targets only: the callers are the native beam searches)
static int server_start(klm_model *m) {
    klm_model::StepServer &sv = m->srv;
    KlmTcPlan &tp = m->tc;
    if (sv.active) return KLM_OK;
    const auto t_s0 = std::chrono::steady_clock::now();
    if (!(tp.ok && tp.df_ok && m->knobs.tc_df && m->knobs.step_mapped && m->knobs.step_server) || tp.n_lse > 8) return KLM_OK;
    int rc = server_ensure_map(m, (size_t)KLM_STEP_MAX * 4 + 256); if (rc) return rc;
    if (!sv.cmd_host) {
        CK(cudaHostAlloc(&sv.cmd_host, sizeof(KlmStepCmd), cudaHostAllocMapped));
        CK(cudaHostGetDevicePointer(reinterpret_cast<void **>(&sv.cmd_host_dev), sv.cmd_host, 0));
        CK(cudaHostAlloc(&sv.done_host, 64, cudaHostAllocMapped));
        CK(cudaHostGetDevicePointer(reinterpret_cast<void **>(&sv.done_host_dev), sv.done_host, 0));
        CK(cudaMalloc(&sv.cmd_dev, sizeof(KlmStepCmd)));
    }
    memset(sv.cmd_host, 0, offsetof(KlmStepCmd, args));
    for (int i = 0; i < 16; ++i) sv.done_host[i] = 0;
    sv.seq = 0;
    *reinterpret_cast<volatile int32_t *>(m->h_map) = 0;
    CK(cudaMemsetAsync(sv.cmd_dev, 0, offsetof(KlmStepCmd, args), m->stream));
    static thread_local KlmTcParams q;
    memset(&q, 0, offsetof(KlmTcParams, step));
    q.W = m->W; q.D = m->D; q.V = m->V; q.Vp = tp.Vp; q.Nc = m->Nc;
    q.B = KLM_TC_STREAMS; q.T_max = 1; q.stages = tp.stages;
    for (int k = 0; k < m->Nc; ++k) { q.ctx_rows[k] = m->cfg.ctx_rows[k]; q.Q[k] = m->Q[k]; }
    q.P = m->P;
    for (int l = 0; l < m->D; ++l) q.bias[l] = m->bias[l];
    q.wimg = tp.d_wimg; q.roles = tp.d_roles;
    q.state = m->state;
    q.slots_in = sv.cmd_dev->args.slots_in; q.slots_out = sv.cmd_dev->args.slots_out;
    q.chars = sv.cmd_dev->args.chars; q.targets = sv.cmd_dev->args.targets; q.ctx = sv.cmd_dev->args.ctx;
    char *db = reinterpret_cast<char *>(m->d_map);
    q.out = reinterpret_cast<float *>(db + 64); q.full_rows = 0;
    q.barrier = m->barrier; q.err = tp.err;
    q.pub_release = m->knobs.tc_pub_release; q.spin_limit = m->knobs.tc_spin_limit; q.fault = m->knobs.tc_fault;
    tc_df_params(m, q);
    q.step_groups = 0; q.n_hyp = 0; q.step_inline = 0;
    q.err_host = reinterpret_cast<int *>(db);
    q.server = 1; q.cmd_host = sv.cmd_host_dev; q.cmd_dev = sv.cmd_dev; q.srv_done = sv.done_host_dev;
    CK(klm_launch_seq_df(q, tp.grid, tp.df_smem, m->stream));
    m->launches++; m->tc_launches++; m->df_launches++;
    sv.active = true; sv.starts++;
    sv.t_start += std::chrono::duration<double>(std::chrono::steady_clock::now() - t_s0).count();
    return KLM_OK;
}

static int server_step(klm_model *m, int n, const int32_t *chars, const int32_t *ctx, const int32_t *slots_in,
                       const int32_t *slots_out, const int32_t *targets, float *out) {
    klm_model::StepServer &sv = m->srv;
    KlmTcPlan &tp = m->tc;
    const volatile float *h_out = reinterpret_cast<const volatile float *>(m->h_map + 64);
    const volatile int32_t *h_err = reinterpret_cast<const volatile int32_t *>(m->h_map);
    for (int s0 = 0; s0 < n; s0 += KLM_STEP_MAX) {
        const auto tm0 = std::chrono::steady_clock::now();
        const int nb = std::min(KLM_STEP_MAX, n - s0);
        KlmStepCmd *c = sv.cmd_host;
        memcpy(c->args.chars, chars + s0, (size_t)nb * 4);
        memcpy(c->args.targets, targets + s0, (size_t)nb * 4);
        memcpy(c->args.slots_in, slots_in + s0, (size_t)nb * 4);
        memcpy(c->args.slots_out, slots_out + s0, (size_t)nb * 4);
        for (int i = 0; i < nb * m->Nc; ++i) c->args.ctx[i] = ctx ? ctx[(size_t)s0 * m->Nc + i] : 0;
        c->n_hyp = nb; c->groups = (nb + KLM_TC_STREAMS - 1) / KLM_TC_STREAMS;
        c->epoch = tp.epoch; c->tick_base = tp.tick_base;
        __atomic_store_n(&c->seq, ++sv.seq, __ATOMIC_RELEASE);
        tp.tick_base += (unsigned)c->groups; tp.epoch += (unsigned)c->groups + 1u;
        const auto tm1 = std::chrono::steady_clock::now();
        // completion: every LSE CTA reports the step number
        unsigned long long spins = 0;
        for (;;) {
            bool all = true;
            for (int i = 0; i < tp.n_lse; ++i)
                if (__atomic_load_n(&sv.done_host[i], __ATOMIC_ACQUIRE) != sv.seq) { all = false; break; }
            if (all) break;
            if ((++spins & 0xFFF) == 0) {
                if (*h_err) break;
                if (std::chrono::duration<double>(std::chrono::steady_clock::now() - tm1).count() > 20.0) {
                    sv.active = false;                        // (the instance leaves by its own watchdog)
                    cudaStreamSynchronize(m->stream);
                    tc_df_reset(m);
                    return fail(KLM_ECUDA, "klm_seq_df (step server): step %u did not complete within 20 s", sv.seq);
                }
            }
        }
        if (*h_err) {
            const int code = *h_err;
            sv.active = false;
            cudaStreamSynchronize(m->stream);
            tc_df_reset(m);
            return fail(KLM_ECUDA, "klm_seq_df (step server): device-side wait timed out (code %d)", code);
        }
        for (int i = 0; i < nb; ++i) out[s0 + i] = h_out[i];
        sv.steps++;
        if (m->timing) {       // device-side stamps of the step (globaltimer, ns): doorbell seen -> device bell rung -> scores out
            const volatile unsigned long long *ts = reinterpret_cast<const volatile unsigned long long *>(sv.done_host + 8);
            sv.t_relay += (double)(ts[1] - ts[0]) * 1e-9; sv.t_device += (double)(ts[2] - ts[1]) * 1e-9;
        }
        if (m->timing) {
            const auto tm2 = std::chrono::steady_clock::now();
            m->t_steps++;
            m->t_step_submit += std::chrono::duration<double>(tm1 - tm0).count();
            m->t_step_wait += std::chrono::duration<double>(tm2 - tm1).count();
            m->t_step_total += std::chrono::duration<double>(tm2 - tm0).count();
        }
    }
    return KLM_OK;
}

static int step_impl(klm_model *m, int32_t n, const int32_t *chars, const int32_t *ctx, const int32_t *slots_in,
                     const int32_t *slots_out, const int32_t *targets, float *out, int mem, int topk, int32_t *out_index) {
    if (!m) return fail(KLM_EINVAL, "NULL model");
    if (n <= 0) return KLM_OK;
    if (!chars || !out) return fail(KLM_EINVAL, "klm_step: NULL argument");
    if (topk && (topk < 1 || topk > m->V || mem != KLM_MEM_HOST || !out_index || targets))
        return fail(KLM_EINVAL, "klm_step_topk: k must be in 1..voc_size, host buffers, no targets");
    if (m->Nc > 0 && !ctx) return fail(KLM_EINVAL, "klm_step: model has %d context variables but ctx is NULL", m->Nc);
    CK(cudaSetDevice(m->device));
    const auto tm0 = std::chrono::steady_clock::now();
    int rc = check_slots(m, slots_in, n, true, "klm_step"); if (rc) return rc;
    rc = check_slots(m, slots_out, n, false, "klm_step"); if (rc) return rc;
    for (int k = 0; k < m->Nc; ++k)
        for (int s = 0; s < n; ++s)
            if (ctx[s * m->Nc + k] < 0 || ctx[s * m->Nc + k] >= m->cfg.ctx_rows[k])
                return fail(KLM_EINVAL, "klm_step: context %d of hypothesis %d is %d, table has %d rows", k, s,
                            ctx[s * m->Nc + k], m->cfg.ctx_rows[k]);
    const int full = targets ? 0 : 1;
    const size_t nout = full ? (size_t)n * m->V : (size_t)n;
    if (m->srv.wanted || m->srv.active) {
        // (steps of <= 8 hypotheses keep the FMA kernel -- the same kernel per step size as without the instance, so that all
        //  forms of the beam search see the same numbers; the instance leaves for them and comes back with the next wide step)
        if (m->srv.wanted && mem == KLM_MEM_HOST && !topk && targets && n > KLM_MODEK_MAXB) {
            if (!m->srv.active) { rc = server_start(m); if (rc) return rc; }
            if (m->srv.active) return server_step(m, n, chars, ctx, slots_in, slots_out, targets, out);
        }
        rc = server_stop(m); if (rc) return rc;
    }
    if (mem == KLM_MEM_HOST && !topk && n > KLM_MODEK_MAXB && n <= KLM_TC_STREAMS * KLM_DF_MAXG && m->tc.ok && m->tc.df_ok &&
        m->knobs.tc_df && m->knobs.step_mapped) {
        // Beam step, zero copy: the per-hypothesis arrays travel INSIDE the kernel parameters (KlmStepArgs), the scores and
        // the watchdog word are written by the kernel straight into MAPPED page-locked host memory.  One launch and one
        // stream synchronisation per step -- no copy-engine transfer in front of or behind the kernel.
        const size_t need = nout * 4 + 256;
        if (need > m->h_map_bytes) {
            CK(cudaStreamSynchronize(m->stream));
            if (m->h_map) cudaFreeHost(m->h_map);
            m->h_map = nullptr; m->h_map_bytes = 0;
            CK(cudaHostAlloc(&m->h_map, need * 2, cudaHostAllocMapped));
            CK(cudaHostGetDevicePointer(&m->d_map, m->h_map, 0));
            m->h_map_bytes = need * 2;
        }
        int32_t *h_err = reinterpret_cast<int32_t *>(m->h_map);
        float *h_out = reinterpret_cast<float *>(m->h_map + 64);
        h_err[0] = 0;
        char *db = reinterpret_cast<char *>(m->d_map);
        const StepHostArgs inl = {chars, targets, slots_in, slots_out, ctx};
        CK(cudaEventRecord(m->ev0, m->stream));
        rc = launch_step_df(m, n, nullptr, nullptr, nullptr, nullptr, nullptr, reinterpret_cast<float *>(db + 64), full,
                            reinterpret_cast<int *>(db), &inl);
        if (rc) return rc;
        CK(cudaEventRecord(m->ev1, m->stream));
        m->timed = true;
        const auto tm1 = std::chrono::steady_clock::now();
        CK(cudaStreamSynchronize(m->stream));
        memcpy(out, h_out, nout * 4);
        if (m->timing) {
            const auto tm2 = std::chrono::steady_clock::now();
            float ms = 0.f;
            cudaEventElapsedTime(&ms, m->ev0, m->ev1);
            m->t_steps++; m->t_step_kernel_ms += ms;
            m->t_step_submit += std::chrono::duration<double>(tm1 - tm0).count();
            m->t_step_wait += std::chrono::duration<double>(tm2 - tm1).count();
            m->t_step_total += std::chrono::duration<double>(std::chrono::steady_clock::now() - tm0).count();
        }
        if (h_err[0]) {
            const int code = h_err[0];
            tc_df_reset(m);
            return fail(KLM_ECUDA, "klm_seq_tc: device-side wait timed out (code %d)", code);
        }
        return KLM_OK;
    }
    rc = ensure_pos_cap(m, (size_t)n, mem == KLM_MEM_HOST ? nout : 0); if (rc) return rc;
    // chars/targets are per-hypothesis host arrays: they ride along as position arrays of length n,
    // through page-locked staging
    const size_t stage_need = (size_t)n * 8 + nout * 4 + 64;
    if (mem == KLM_MEM_HOST && stage_need > m->h_step_bytes) {
        CK(cudaStreamSynchronize(m->stream));
        if (m->h_step) cudaFreeHost(m->h_step);
        m->h_step = nullptr; m->h_step_bytes = 0;
        CK(cudaHostAlloc(&m->h_step, stage_need * 2, cudaHostAllocDefault));
        m->h_step_bytes = stage_need * 2;
    }
    const int32_t *src_chars = chars, *src_targets = targets;
    if (mem == KLM_MEM_HOST) {
        int32_t *hs = reinterpret_cast<int32_t *>(m->h_step);
        memcpy(hs, chars, (size_t)n * 4);
        if (targets) memcpy(hs + n, targets, (size_t)n * 4);
        src_chars = hs; src_targets = hs + n;
    }
    CK(cudaMemcpyAsync(m->d_chars, src_chars, (size_t)n * 4, cudaMemcpyDefault, m->stream));       // (host or device, by `mem`)
    if (targets) CK(cudaMemcpyAsync(m->d_targets, src_targets, (size_t)n * 4, cudaMemcpyDefault, m->stream));
    float *dout = (mem == KLM_MEM_HOST) ? m->d_out : out;
    CK(cudaEventRecord(m->ev0, m->stream));
    if (n > KLM_MODEK_MAXB && m->tc.ok && m->tc.df_ok && m->knobs.tc_df) {
        // beams: the dataflow kernel's step mode, up to 256 hypotheses per launch
        const int bmax = KLM_TC_STREAMS * KLM_DF_MAXG;
        for (int s0 = 0; s0 < n; s0 += bmax) {
            const int nb = std::min(bmax, n - s0);
            rc = run_step_df(m, nb, slots_in + s0, slots_out + s0, ctx ? ctx + (size_t)s0 * m->Nc : nullptr, m->d_chars + s0,
                             targets ? m->d_targets + s0 : nullptr, dout + (full ? (size_t)s0 * m->V : (size_t)s0), full);
            if (rc) return rc;
        }
    } else {
        const int bmax = batch_max(m, n);
        std::vector<int64_t> off0;
        for (int s0 = 0; s0 < n; s0 += bmax) {
            const int nb = std::min(bmax, n - s0);
            off0.resize(nb + 1);
            for (int i = 0; i <= nb; ++i) off0[i] = s0 + i;
            rc = run_batch(m, nb, slots_in + s0, slots_out + s0, off0.data(), nullptr, ctx ? ctx + (size_t)s0 * m->Nc : nullptr,
                           m->d_chars, targets ? m->d_targets : nullptr, dout, full);
            if (rc) return rc;
        }
    }
    CK(cudaEventRecord(m->ev1, m->stream));
    m->timed = true;
    if (mem == KLM_MEM_HOST) {
        float *hout = reinterpret_cast<float *>(m->h_step + (size_t)n * 8);
        size_t ncopy = nout;
        if (topk) {
            // the rows stay on the device: k (index, probability) pairs per hypothesis come back instead of voc_size floats
            if ((size_t)n * topk * 8 > m->topk_bytes) {
                CK(cudaStreamSynchronize(m->stream));
                cudaFree(m->d_topk); m->d_topk = nullptr; m->topk_bytes = 0;
                CK(cudaMalloc(&m->d_topk, (size_t)n * topk * 8 * 2));
                m->topk_bytes = (size_t)n * topk * 8 * 2;
            }
            int *d_idx = reinterpret_cast<int *>(m->d_topk);
            float *d_prob = reinterpret_cast<float *>(m->d_topk) + (size_t)n * topk;
            klm_topk_rows<<<(n + 3) / 4, 128, 0, m->stream>>>(m->d_out, n, m->V, topk, d_idx, d_prob);
            m->launches++;
            ncopy = (size_t)n * topk * 2;                              // indices, then probabilities
            CK(cudaMemcpyAsync(hout, m->d_topk, ncopy * 4, cudaMemcpyDeviceToHost, m->stream));
        } else {
            CK(cudaMemcpyAsync(hout, m->d_out, nout * 4, cudaMemcpyDeviceToHost, m->stream));
        }
        rc = fetch_errors_async(m); if (rc) return rc;
        const auto tm1 = std::chrono::steady_clock::now();
        CK(cudaStreamSynchronize(m->stream));
        if (topk) {
            memcpy(out_index, hout, (size_t)n * topk * 4);
            memcpy(out, hout + (size_t)n * topk, (size_t)n * topk * 4);
        } else {
            memcpy(out, hout, nout * 4);
        }
        if (m->timing) {
            const auto tm2 = std::chrono::steady_clock::now();
            float ms = 0.f;
            cudaEventElapsedTime(&ms, m->ev0, m->ev1);
            m->t_steps++; m->t_step_kernel_ms += ms;
            m->t_step_submit += std::chrono::duration<double>(tm1 - tm0).count();
            m->t_step_wait += std::chrono::duration<double>(tm2 - tm1).count();
            m->t_step_total += std::chrono::duration<double>(std::chrono::steady_clock::now() - tm0).count();
        }
        return check_fetched_errors(m);
    }
    return KLM_OK;
}

extern "C" int klm_step(klm_model *m, int32_t n, const int32_t *chars, const int32_t *ctx, const int32_t *slots_in,
                        const int32_t *slots_out, const int32_t *targets, float *out, int mem) {
    return step_impl(m, n, chars, ctx, slots_in, slots_out, targets, out, mem, 0, nullptr);
}

extern "C" int klm_step_topk(klm_model *m, int32_t n, const int32_t *chars, const int32_t *ctx, const int32_t *slots_in,
                             const int32_t *slots_out, int32_t k, int32_t *out_index, float *out_prob) {
    return step_impl(m, n, chars, ctx, slots_in, slots_out, nullptr, out_prob, KLM_MEM_HOST, k, out_index);
}

extern "C" int klm_get_info(klm_model *m, klm_info *out) {
    if (!m || !out) return fail(KLM_EINVAL, "klm_get_info: NULL argument");
    memset(out, 0, sizeof *out);
    out->sm_count = m->sm_count;
    out->grid = m->plan_b.grid;
    out->weights_in_smem = m->plan_b.wsmem;
    out->smem_bytes = (int32_t)m->plan_b.smem;
    out->kernel_launches = m->launches;
    out->max_streams = m->tc.ok ? ((m->tc.wide_ok && m->knobs.tc_wide) ? KLM_TW_MAX_STREAMS : KLM_TC_STREAMS) : m->plan_b.max_streams;
    out->tc_enabled = m->tc.ok ? (m->tc.stream_b ? 3 : 1) : 0; // 3: streamed-weights form (images stay in global memory / L2)
    out->tc_grid = m->tc.grid;
    out->tc_smem_bytes = (int32_t)m->tc.smem;
    out->tc_launches = m->tc_launches;
    out->cl_launches = m->cl_launches; out->cl_cluster = m->cl.C;
    out->tc_dataflow = (m->tc.ok && m->tc.df_ok && m->knobs.tc_df) ? 1 : 0;
    out->tc_wide = (m->tc.ok && m->tc.wide_ok && m->knobs.tc_wide) ? 1 : 0;
    out->tc_wide_launches = m->tc.wide_launches;
    if (m->timed) {
        CK(cudaSetDevice(m->device));
        CK(cudaEventSynchronize(m->ev1));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, m->ev0, m->ev1));
        out->last_kernel_ms = ms;
    }
    return KLM_OK;
}

// ------------------------------------------------------------------------------------------
// Self-test of the tcgen05 / TMEM / bulk-copy building blocks (tests/): D[128 x N] = A . B^T for fp16
// operand images in the kernel's shared-memory layout ([K/8][rows][8] halves).  Host pointers.
extern "C" int klm_selftest_umma(int device, int32_t N, int32_t K, const uint16_t *A_img, const uint16_t *B_img, float *D) {
    if (!A_img || !B_img || !D) return fail(KLM_EINVAL, "klm_selftest_umma: NULL argument");
    if (N < 16 || N > 256 || N % 16 || K < 16 || K % 16 || (size_t)K * (128 + N) * 2 > 200 * 1024)
        return fail(KLM_EINVAL, "klm_selftest_umma: unsupported shape N=%d K=%d", N, K);
    CK(cudaSetDevice(device));
    __half *dA = nullptr, *dB = nullptr; float *dD = nullptr; int *derr = nullptr;
    CK(cudaMalloc(&dA, (size_t)K * 128 * 2)); CK(cudaMalloc(&dB, (size_t)K * N * 2));
    CK(cudaMalloc(&dD, (size_t)128 * N * 4)); CK(cudaMalloc(&derr, 4));
    CK(cudaMemcpy(dA, A_img, (size_t)K * 128 * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, B_img, (size_t)K * N * 2, cudaMemcpyHostToDevice));
    CK(cudaMemset(dD, 0, (size_t)128 * N * 4)); CK(cudaMemset(derr, 0, 4));
    CK(klm_launch_umma_selftest(dA, dB, dD, N, K, derr, 0));
    CK(cudaDeviceSynchronize());
    int code = 0;
    CK(cudaMemcpy(&code, derr, 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(D, dD, (size_t)128 * N * 4, cudaMemcpyDeviceToHost));
    cudaFree(dA); cudaFree(dB); cudaFree(dD); cudaFree(derr);
    if (code) return fail(KLM_ECUDA, "klm_selftest_umma: device-side wait timed out (code %d)", code);
    return KLM_OK;
}

// Debug timeline of the tensor-core kernel.  out == NULL: arm tracing of KLM_TC_TRACE_TICKS ticks from
// `first_tick` for subsequent launches.  out != NULL: copy the [grid][ticks][8] SM-clock samples and the
// role table ([grid][4] ints: kind, layer, npairs, sigma) to host.
extern "C" int klm_tc_trace(klm_model *m, int32_t first_tick, uint64_t *out, int32_t *roles_out) {
    if (!m) return fail(KLM_EINVAL, "NULL model");
    if (!m->tc.ok) return fail(KLM_EINVAL, "klm_tc_trace: the tensor-core plan is not enabled for this model");
    CK(cudaSetDevice(m->device));
    KlmTcPlan &tp = m->tc;
    const size_t n = (size_t)tp.grid * KLM_TC_TRACE_TICKS * 8;
    if (!out) {
        if (!tp.trace) CK(cudaMalloc(&tp.trace, n * 8));
        CK(cudaMemset(tp.trace, 0, n * 8));
        tp.trace_t0 = first_tick;
        return KLM_OK;
    }
    if (!tp.trace) return fail(KLM_EINVAL, "klm_tc_trace: not armed");
    CK(cudaStreamSynchronize(m->stream));
    CK(cudaMemcpy(out, tp.trace, n * 8, cudaMemcpyDeviceToHost));
    if (roles_out)
        for (int i = 0; i < tp.grid; ++i) {
            roles_out[4 * i] = tp.roles[i].kind; roles_out[4 * i + 1] = tp.roles[i].layer;
            roles_out[4 * i + 2] = tp.roles[i].npairs; roles_out[4 * i + 3] = tp.roles[i].sigma;
        }
    return KLM_OK;
}

// ------------------------------------------------------------------------------------------
// One lattice edge of the beam search, natively: replaces the inner loop of Rater.rate_best
// (rating.py:789-852) for one edge -- candidate = (incoming hypothesis, alternative), character-
// synchronous batched steps, worst-first pops, projected-cost ordering with Python's
// bisect.insort_left (replicated comparison by comparison, including its behaviour on the not yet
// sorted initial list), the +15 bit cardinality break, the +2.5 bit threshold and the truncation of
// the waiting list.  Clustering (beam_clustering_dist) and joins into an already populated node are
// left to the host-language path.
namespace {
struct BeamCand { int hyp, alt, len, alt_len, slot; bool own; double cum, conf_cost; };
struct BeamLists {
    std::vector<BeamCand> c;
    double pro(int i) const { return c[i].cum + 0.5 * (double)(c[i].alt_len - c[i].len); }
    void insort_left(std::vector<int> &a, int x) const {       // bisect.insort_left with Node.__lt__ (:1300-1318)
        size_t lo = 0, hi = a.size();
        const double px = pro(x);
        while (lo < hi) {
            const size_t mid = (lo + hi) / 2;
            if (pro(a[mid]) < px) lo = mid + 1; else hi = mid;
        }
        a.insert(a.begin() + lo, x);
    }
};
}  // namespace

// One edge: candidates = hypotheses x alternatives; `done` receives the finished candidates in beam order (indices into
// L.c); state slots come from / go back to `free_slots`.
static int beam_edge_core(klm_model *m, BeamLists &L, int n_hyp, const int32_t *hyp_slot, const int32_t *hyp_last_char,
                          const double *hyp_cum_cost, int n_alt, const int32_t *alt_offsets, const int32_t *alt_chars,
                          const double *alt_conf, const int32_t *ctx, double lm_weight, int batch_size,
                          std::vector<int> &free_slots, std::vector<int> &done, int64_t *n_steps, int64_t *n_hyp_steps) {
    std::vector<int> waiting, batch;
    const double log2d = std::log(2.0);
    int max_len = 0;
    for (int a = 0; a < n_alt; ++a) max_len = std::max(max_len, alt_offsets[a + 1] - alt_offsets[a]);
    L.c.clear();
    done.clear();
    for (int h = 0; h < n_hyp; ++h)
        for (int a = 0; a < n_alt; ++a) {                        // beam-major, not re-sorted (:789-795)
            BeamCand c;
            c.hyp = h; c.alt = a; c.len = 0; c.alt_len = alt_offsets[a + 1] - alt_offsets[a];
            c.slot = hyp_slot[h]; c.own = false; c.cum = hyp_cum_cost[h];
            c.conf_cost = -(std::log(std::max(alt_conf[a], 1e-99)) / log2d) * (1.0 - lm_weight);
            waiting.push_back((int)L.c.size());
            L.c.push_back(c);
        }
    auto release = [&](int i) { if (L.c[i].own) { free_slots.push_back(L.c[i].slot); L.c[i].own = false; } };
    const int rounds = 3 * max_len;                              // :797
    std::vector<int32_t> feed, want, s_in, s_out, cx;
    std::vector<float> logp;
    for (int r = 0; r < rounds; ++r) {
        batch.clear();
        while (!waiting.empty()) {                               // most expensive first (:800-810)
            const int i = waiting.back();
            waiting.pop_back();
            if (L.c[i].len == L.c[i].alt_len) L.insort_left(done, i);
            else L.insort_left(batch, i);
            if ((int)batch.size() >= batch_size) break;
        }
        if (batch.empty()) break;
        if (!done.empty() && L.c[batch[0]].cum >= L.c[done[0]].cum + 15.0) {   // :816-818
            for (int i : batch) release(i);
            for (int i : waiting) release(i);
            batch.clear(); waiting.clear();
            break;
        }
        const int n = (int)batch.size();
        if ((int)free_slots.size() < n) return fail(KLM_ESTATE, "klm_beam_edge: out of scratch state slots");
        feed.resize(n); want.resize(n); s_in.resize(n); s_out.resize(n); logp.resize(n);
        cx.resize((size_t)n * std::max(1, m->Nc));
        for (int k = 0; k < n; ++k) {
            const BeamCand &c = L.c[batch[k]];
            feed[k] = c.len ? alt_chars[alt_offsets[c.alt] + c.len - 1] : hyp_last_char[c.hyp];   // :821-824
            want[k] = alt_chars[alt_offsets[c.alt] + c.len];
            s_in[k] = c.slot;
            s_out[k] = free_slots.back(); free_slots.pop_back();
            for (int j = 0; j < m->Nc; ++j) cx[(size_t)k * m->Nc + j] = ctx ? ctx[j] : 0;
        }
        int rc = klm_step(m, n, feed.data(), m->Nc ? cx.data() : nullptr, s_in.data(), s_out.data(), want.data(), logp.data(),
                          KLM_MEM_HOST);
        if (rc) return rc;
        if (n_steps) ++*n_steps;
        if (n_hyp_steps) *n_hyp_steps += n;
        for (int k = 0; k < n; ++k) {                            // ascending projected cost (:827-849)
            const int i = batch[k];
            BeamCand &c = L.c[i];
            const float prob = expf(logp[k]);                    // float32, like the softmax output it replaces
            c.cum += -(std::log(std::max((double)prob, 1e-99)) / log2d) * lm_weight + c.conf_cost;
            c.len += 1;
            release(i);
            c.slot = s_out[k]; c.own = true;
            if (!waiting.empty() && c.cum >= L.c[waiting[0]].cum + 2.5) { release(i); continue; }   // :847-848
            L.insort_left(waiting, i);
        }
        const size_t keep = (size_t)rounds * batch_size;         // :851
        if (waiting.size() > keep) {
            for (size_t k = keep; k < waiting.size(); ++k) release(waiting[k]);
            waiting.resize(keep);
        }
    }
    // (candidates still waiting when the rounds are used up keep their slots until the caller drops the edge)
    for (int i : waiting) release(i);
    return KLM_OK;
}

// the resident step instance lives exactly as long as one native beam-search call
struct ServerGuard {
    klm_model *m;
    int rc;
    explicit ServerGuard(klm_model *m_) : m(m_), rc(KLM_OK) { m->srv.wanted = true; }      // (started by the first wide step)
    int stop() { m->srv.wanted = false; return server_stop(m); }
    ~ServerGuard() { m->srv.wanted = false; server_stop(m); }
};

extern "C" int klm_beam_edge(klm_model *m, int32_t n_hyp, const int32_t *hyp_slot, const int32_t *hyp_last_char,
                             const double *hyp_cum_cost, int32_t n_alt, const int32_t *alt_offsets,
                             const int32_t *alt_chars, const double *alt_conf, const int32_t *ctx, double lm_weight,
                             int32_t batch_size, int32_t n_scratch, const int32_t *scratch_slots, int32_t max_out,
                             int32_t *out_hyp, int32_t *out_alt, double *out_cum_cost, int32_t *out_slot, int32_t *n_out,
                             int64_t *n_steps, int64_t *n_hyp_steps) {
    if (!m || !hyp_slot || !hyp_last_char || !hyp_cum_cost || !alt_offsets || !alt_chars || !alt_conf || !scratch_slots ||
        !out_hyp || !out_alt || !out_cum_cost || !out_slot || !n_out)
        return fail(KLM_EINVAL, "klm_beam_edge: NULL argument");
    if (n_hyp < 0 || n_alt < 1 || batch_size < 1 || max_out < 0) return fail(KLM_EINVAL, "klm_beam_edge: bad sizes");
    *n_out = 0;
    const auto te0 = std::chrono::steady_clock::now();
    struct EdgeTimer { klm_model *m; std::chrono::steady_clock::time_point t0;
        ~EdgeTimer() { if (m->timing) { m->t_edges++; m->t_edge_total += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); } } } edge_timer{m, te0};
    if (n_steps) *n_steps = 0;
    if (n_hyp_steps) *n_hyp_steps = 0;
    BeamLists L;
    std::vector<int> done, free_slots(scratch_slots, scratch_slots + n_scratch);
    CK(cudaSetDevice(m->device));
    ServerGuard server(m);
    if (server.rc) return server.rc;
    int rc = beam_edge_core(m, L, n_hyp, hyp_slot, hyp_last_char, hyp_cum_cost, n_alt, alt_offsets, alt_chars, alt_conf, ctx,
                            lm_weight, batch_size, free_slots, done, n_steps, n_hyp_steps);
    if (rc) return rc;
    rc = server.stop(); if (rc) return rc;
    const int no = std::min<int>(max_out, (int)done.size());
    for (int k = 0; k < no; ++k) {
        const BeamCand &c = L.c[done[k]];
        out_hyp[k] = c.hyp; out_alt[k] = c.alt; out_cum_cost[k] = c.cum; out_slot[k] = c.own ? c.slot : -2;   // -2: the parent's state
    }
    *n_out = no;
    return KLM_OK;
}

// A whole LINEAR lattice (a chain of edges: what wrapper/rate.py builds from a PAGE, :343-424) in one call: the
// per-edge search above for edge after edge, the first `beam_width` finished candidates of an edge being the incoming
// hypotheses of the next one (rating.py:763-852 for a graph without branches).  The host language gets a table of
// back-pointers and builds tree nodes only for what it looks at afterwards.
//
//   n_edges, edge_alt_first [n_edges+1]          alternatives edge_alt_first[e] .. edge_alt_first[e+1]-1 belong to edge e
//   alt_offsets [n_alts+1], alt_chars, alt_conf  characters (indices) and confidences of all alternatives
//   out_first [n_edges+1]                        kept candidates of edge e are rows out_first[e] .. out_first[e+1]-1 of
//   out_hyp / out_alt / out_cum_cost [max_out]   (index into the kept rows of the previous edge -- or the incoming beam for
//                                                edge 0 --, alternative index relative to the edge, cumulative cost)
//   final_slot [beam_width]                      state slots of the last edge's candidates (scratch slots now owned by the
//                                                caller, or -2 - i: the state of incoming hypothesis i, untouched)
// scratch_slots: at least (beam_width + batch_size) * (largest number of alternatives of an edge + 2) free slots.
extern "C" int klm_beam_lattice(klm_model *m, int32_t n_hyp, const int32_t *hyp_slot, const int32_t *hyp_last_char,
                                const double *hyp_cum_cost, int32_t n_edges, const int32_t *edge_alt_first,
                                const int32_t *alt_offsets, const int32_t *alt_chars, const double *alt_conf,
                                const int32_t *ctx, double lm_weight, int32_t batch_size, int32_t beam_width,
                                int32_t n_scratch, const int32_t *scratch_slots, int32_t max_out, int32_t *out_first,
                                int32_t *out_hyp, int32_t *out_alt, double *out_cum_cost, int32_t *final_slot,
                                int64_t *n_steps, int64_t *n_hyp_steps) {
    if (!m || !hyp_slot || !hyp_last_char || !hyp_cum_cost || !edge_alt_first || !alt_offsets || !alt_chars || !alt_conf ||
        !scratch_slots || !out_first || !out_hyp || !out_alt || !out_cum_cost || !final_slot)
        return fail(KLM_EINVAL, "klm_beam_lattice: NULL argument");
    if (n_hyp < 0 || n_edges < 0 || batch_size < 1 || beam_width < 1 || max_out < 0) return fail(KLM_EINVAL, "klm_beam_lattice: bad sizes");
    if (n_steps) *n_steps = 0;
    if (n_hyp_steps) *n_hyp_steps = 0;
    BeamLists L;
    std::vector<int> done, free_slots(scratch_slots, scratch_slots + n_scratch);
    // the current beam: state slot (negative = not ours: -1 zero state, <= -2 incoming hypothesis -2-i), whether the slot is one
    // of ours, how many rows of the beam share it, last character, cost
    std::vector<int32_t> b_slot(hyp_slot, hyp_slot + n_hyp), b_last(hyp_last_char, hyp_last_char + n_hyp), b_origin(n_hyp);
    std::vector<double> b_cum(hyp_cum_cost, hyp_cum_cost + n_hyp);
    std::vector<char> b_own(n_hyp, 0);
    for (int i = 0; i < n_hyp; ++i) b_origin[i] = -2 - i;
    out_first[0] = 0;
    int nout = 0;
    CK(cudaSetDevice(m->device));
    ServerGuard server(m);
    if (server.rc) return server.rc;
    for (int e = 0; e < n_edges; ++e) {
        const int a0 = edge_alt_first[e], n_alt = edge_alt_first[e + 1] - a0;
        if (n_alt < 1) return fail(KLM_EINVAL, "klm_beam_lattice: edge %d has no alternatives", e);
        // offsets of this edge's alternatives, relative to alt_chars
        int rc = beam_edge_core(m, L, (int)b_slot.size(), b_slot.data(), b_last.data(), b_cum.data(), n_alt, alt_offsets + a0,
                                alt_chars, alt_conf + a0, ctx, lm_weight, batch_size, free_slots, done, n_steps, n_hyp_steps);
        if (rc) return rc;
        const int keep = std::min<int>(beam_width, (int)done.size());
        if (nout + keep > max_out) return fail(KLM_EINVAL, "klm_beam_lattice: output arrays too small");
        std::vector<int32_t> n_slot(keep), n_last(keep), n_origin(keep);
        std::vector<double> n_cum(keep);
        std::vector<char> n_own(keep), parent_kept(b_slot.size(), 0);
        for (int k = 0; k < keep; ++k) {
            const BeamCand &c = L.c[done[k]];
            out_hyp[nout + k] = c.hyp; out_alt[nout + k] = c.alt; out_cum_cost[nout + k] = c.cum;
            n_cum[k] = c.cum;
            n_last[k] = c.alt_len ? alt_chars[alt_offsets[a0 + c.alt] + c.alt_len - 1] : b_last[c.hyp];
            if (c.own) { n_slot[k] = c.slot; n_own[k] = 1; n_origin[k] = c.slot; }
            else {       // an empty alternative: the candidate lives on with its parent's state
                n_slot[k] = b_slot[c.hyp]; n_own[k] = b_own[c.hyp]; n_origin[k] = b_origin[c.hyp]; parent_kept[c.hyp] = 1;
            }
        }
        for (size_t k = keep; k < done.size(); ++k)               // finished, but outside the beam
            if (L.c[done[k]].own) free_slots.push_back(L.c[done[k]].slot);
        for (size_t i = 0; i < b_slot.size(); ++i)                // the previous beam's states are no longer needed
            if (b_own[i] && !parent_kept[i]) {
                bool dup = false;                                  // (rows of a beam may share a slot through empty alternatives)
                for (size_t j = 0; j < i; ++j) if (b_own[j] && b_slot[j] == b_slot[i] && !parent_kept[j]) dup = true;
                bool still = false;
                for (int k = 0; k < keep; ++k) if (n_own[k] && n_slot[k] == b_slot[i]) still = true;
                if (!dup && !still) free_slots.push_back(b_slot[i]);
            }
        b_slot.swap(n_slot); b_last.swap(n_last); b_cum.swap(n_cum); b_own.swap(n_own); b_origin.swap(n_origin);
        nout += keep;
        out_first[e + 1] = nout;
    }
    for (int k = 0; k < beam_width; ++k) final_slot[k] = k < (int)b_slot.size() ? (b_own[k] ? b_slot[k] : b_origin[k]) : -1;
    return server.stop();
}

// Micro-benchmark behind the roofline of the small-batch (FP32 FMA) kernels: 8 independent FFMA chains per thread, 4 CTAs of
// 256 threads per SM, ~0.5 ms; returns the measured FP32 FMA throughput of the device in TFLOP/s (2 flops per FMA).
__global__ void klm_fma_peak_kernel(float *out, int iters) {
    float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1.f, a2 = a0 + 2.f, a3 = a0 + 3.f, a4 = a0 + 4.f, a5 = a0 + 5.f, a6 = a0 + 6.f, a7 = a0 + 7.f;
    const float b = 1.0000001f, c = 1e-7f;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            a0 = fmaf(a0, b, c); a1 = fmaf(a1, b, c); a2 = fmaf(a2, b, c); a3 = fmaf(a3, b, c);
            a4 = fmaf(a4, b, c); a5 = fmaf(a5, b, c); a6 = fmaf(a6, b, c); a7 = fmaf(a7, b, c);
        }
    }
    const float r = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
    if (r == 123.456f) out[0] = r;                           // (never true: keeps the chains alive)
}
extern "C" int klm_lpt_partition(const int64_t *lengths, int64_t n, int32_t world, int32_t *rank_of) {
    if (n < 0 || world < 1 || (n > 0 && (!lengths || !rank_of))) return fail(KLM_EINVAL, "klm_lpt_partition: bad argument");
    std::vector<int64_t> order((size_t)n);
    for (int64_t i = 0; i < n; ++i) order[(size_t)i] = i;
    std::stable_sort(order.begin(), order.end(), [&](int64_t a, int64_t b) { return lengths[a] > lengths[b]; });
    typedef std::pair<int64_t, int32_t> Load;                              // (load, rank): least loaded first, ties by rank
    std::priority_queue<Load, std::vector<Load>, std::greater<Load>> heap;
    for (int32_t r = 0; r < world; ++r) heap.push(Load(0, r));
    for (int64_t k = 0; k < n; ++k) {
        const int64_t i = order[(size_t)k];
        Load top = heap.top();
        heap.pop();
        rank_of[i] = top.second;
        top.first += lengths[i] + 1;
        heap.push(top);
    }
    return KLM_OK;
}

extern "C" int klm_measure_fma_peak(int device, double *tflops) {
    if (!tflops) return fail(KLM_EINVAL, "klm_measure_fma_peak: NULL argument");
    CK(cudaSetDevice(device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    float *d = nullptr;
    CK(cudaMalloc(&d, 4));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const int grid = prop.multiProcessorCount * 8, threads = 256, iters = 4096;
    double best = 0;
    for (int rep = 0; rep < 4; ++rep) {
        CK(cudaEventRecord(e0, 0));
        klm_fma_peak_kernel<<<grid, threads>>>(d, iters);
        CK(cudaEventRecord(e1, 0));
        CK(cudaEventSynchronize(e1));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        const double fl = 2.0 * 8 * 16 * (double)iters * threads * grid;
        if (rep > 0) best = std::max(best, fl / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d);
    *tflops = best;
    return KLM_OK;
}

// Self-test of the CTA-pair MMA (cta_group::2): A_img = two [K/8][64][8] halves (rows 0..63, 64..127), B_img = two
// [K/8][N/2][8] halves; D receives the raw TMEM dump [2 CTAs][128 lanes][N/2 columns].  Host pointers.
extern "C" int klm_selftest_umma2(int device, int32_t N, int32_t K, const uint16_t *A_img, const uint16_t *B_img, float *D) {
    if (!A_img || !B_img || !D) return fail(KLM_EINVAL, "klm_selftest_umma2: NULL argument");
    if (N < 32 || N > 256 || N % 32 || K < 16 || K % 16 || (size_t)K * (64 + N / 2) * 2 > 200 * 1024)
        return fail(KLM_EINVAL, "klm_selftest_umma2: unsupported shape N=%d K=%d", N, K);
    CK(cudaSetDevice(device));
    __half *dA = nullptr, *dB = nullptr; float *dD = nullptr; int *derr = nullptr;
    CK(cudaMalloc(&dA, (size_t)K * 128 * 2)); CK(cudaMalloc(&dB, (size_t)K * N * 2));
    CK(cudaMalloc(&dD, (size_t)2 * 128 * (N / 2) * 4)); CK(cudaMalloc(&derr, 4));
    CK(cudaMemcpy(dA, A_img, (size_t)K * 128 * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, B_img, (size_t)K * N * 2, cudaMemcpyHostToDevice));
    CK(cudaMemset(dD, 0, (size_t)2 * 128 * (N / 2) * 4)); CK(cudaMemset(derr, 0, 4));
    CK(klm_launch_umma2_selftest(dA, dB, dD, N, K, derr, 0));
    CK(cudaDeviceSynchronize());
    int code = 0;
    CK(cudaMemcpy(&code, derr, 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(D, dD, (size_t)2 * 128 * (N / 2) * 4, cudaMemcpyDeviceToHost));
    cudaFree(dA); cudaFree(dB); cudaFree(dD); cudaFree(derr);
    if (code) return fail(KLM_ECUDA, "klm_selftest_umma2: device-side wait timed out (code %d)", code);
    return KLM_OK;
}
