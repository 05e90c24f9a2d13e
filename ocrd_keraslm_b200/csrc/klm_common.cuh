// Shared definitions of libkeraslm_b200 (device + host).
//
// Math being computed: SURVEY.md Appendix A, i.e. the graph of Rater.configure
// (/root/reference/ocrd_keraslm/lib/rating.py:61-179):
//   u = concat(E[x], C_1[z_1], ...)                      :103-115,125
//   per layer: a = in.K + h.R + b ; i,f,g,o = s(a0),s(a1),tanh(a2),s(a3)   :126-145 (gate order i|f|c|o)
//              c = f*c + i*g ; h = o*tanh(c)
//   p = softmax(E . h_D)                                  :155-168 (tied, no bias)
//
// Device-resident forms (built once in klm_create):
//   P  [V][W] float4      P[x][u] = (E[x] . K_1[:W, g*W+u] + b_1[g*W+u]) for g = i,f,g,o.  The input
//                         of layer 1 is always an embedding row, so its input contraction is a table.
//   Q_k[rows][W] float4   Q_k[z][u] = C_k[z] . K_1[W+off_k : W+off_k+d_k, g*W+u]   (context part)
//   Wl[l][W][K_l] float4  per unit u and input index k the four gate weights; K_1 = W (R_1 only),
//                         K_l = 2W for l > 1 (rows of K_l, then rows of R_l)
//   bias[l][W] float4     l > 1 (layer 1's bias is folded into P)
//   Ep [V4/4][W] float4   output projection: four vocabulary rows of E per "unit"
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/keraslm_b200.h"

// tanh(x) as a quotient num / den with e = exp(-2x): num = 1 - e, den = 1 + e.  For small |x| the numerator cancels
// (e ~ 1: an absolute error of ~1e-7 in a value of ~2|x|; with Keras-default initialisation cell values are ~1e-3 and
// the states drifted 1.7e-4 relative, tests/test_gpu_long.py), so below 1/8 the odd polynomial
// x - x^3/3 + 2x^5/15 - 17x^7/315 takes its place (truncation error 1e-9 relative).  The callers fold the quotient into
// the division they do anyway (sigmoid(i) * tanh(g) = num / ((1 + e_i) * den)).
__device__ __forceinline__ void klm_tanh_parts(float x, float e, float &num, float &den) {
    const float x2 = x * x;
    const float poly = x * fmaf(x2, fmaf(x2, fmaf(x2, -17.f / 315.f, 2.f / 15.f), -1.f / 3.f), 1.f);
    const bool small = fabsf(x) < 0.125f;
    num = small ? poly : 1.f - e;
    den = small ? 1.f : 1.f + e;
}

#define KLM_THREADS      512
#define KLM_WARPS        (KLM_THREADS / 32)
#define KLM_MAX_ITEMS    8      // (unit, 32-stream group) items per warp in the stream-parallel mode
#define KLM_MAX_UNITS_K  4      // units per warp in the K-parallel mode
#define KLM_MODEK_MAXB   8      // K-parallel mode takes up to this many streams
#define KLM_LL_RING      32     // ticks of exchange words kept by the barrier-free K-parallel kernel

#define KLM_ROLE_LSTM 0
#define KLM_ROLE_PROJ 1
#define KLM_ROLE_LSE  2
#define KLM_ROLE_IDLE 3

struct KlmRole {
    int kind;    // KLM_ROLE_*
    int layer;   // LSTM layer (0-based)
    int first;   // first unit (LSTM) / first group of 4 vocabulary rows (PROJ)
    int count;   // number of units / groups owned by this CTA
};

struct KlmSeqParams {
    int W, D, V, V4, Nc;
    int B, Bp, T_max;
    int ctx_rows[KLM_MAX_CTX];
    const float4 *P;
    const float4 *Q[KLM_MAX_CTX];
    const float4 *Wl[KLM_MAX_DEPTH];
    const float4 *bias[KLM_MAX_DEPTH];
    const float4 *Ep;
    float *state;                 // pool [slot][D][2][W]
    const int *slots_in;          // [B] device, -1 = zeros
    const int *slots_out;         // [B] device
    const long long *offsets;     // [B+1] device
    const int *chars;             // device, indexed by absolute position
    const int *targets;           // device or NULL (full rows)
    const int *n_real;            // [B] device or NULL
    const int *ctx;               // [B][Nc] device
    float *out;                   // logp per position, or [pos][V] probabilities
    int full_rows;
    float *hbuf;                  // [D][2][W][Bp]
    float *logits;                // [2][V4][Bp]
    unsigned int *barrier;        // zeroed before launch
    const KlmRole *roles;         // [grid]
    int wsmem;                    // weights slice cached in shared memory
    int stage_floats;             // floats of staging area
    // barrier-free K-parallel kernel (klm_seq_fma_ll): exchanged values as {value, tick} words in rings of KLM_LL_RING ticks
    uint2 *hbuf_ll;               // [KLM_LL_RING][D][W][KLM_MODEK_MAXB]; NULL = use the barrier kernels
    uint2 *logits_ll;             // [KLM_LL_RING][4*V4][KLM_MODEK_MAXB]
    unsigned int *progress;       // [grid] ticks completed per CTA (zeroed before launch)
    int *err;                     // device-side watchdog
    int debug;                    // KLM_LL_DEBUG: every CTA prints its mean clocks per tick spent waiting / computing
};

// One-cluster kernel (klm_seq_cl): geometry chosen by the host
#define KLM_CL_RING   4      // ticks of exchange words kept in every CTA's shared memory
#define KLM_CL_MAXT   4      // tasks (units / groups of 4 vocabulary rows) per warp
#define KLM_CL_PUBS   8      // words a thread publishes per tick at most ((D U C + 4 PU) B <= 8 * 512)
struct KlmClGeom {
    int C;         // CTAs in the cluster (8 or 16)
    int U;         // units of every layer owned by one CTA (W / C)
    int PU;        // groups of 4 vocabulary rows owned by one CTA (ceil(V4 / C))
    int ntask;     // D * U + PU
    int w4;        // float4s of one CTA's weight slices
};

// ---------------------------------------------------------------------------------------------
// Tensor-core sequence kernel (klm_seq_tc.cu)
#define KLM_TC_THREADS      320                    // producer warp, MMA warp, 8 epilogue warps
#define KLM_TC_STREAMS      64                     // accumulator rows 0..63 = hi part, 64..127 = lo part
#define KLM_TC_KC           64                     // k per ring chunk
#define KLM_TC_CHUNK_BYTES  (KLM_TC_KC / 8 * 128 * 16)
#define KLM_TC_OWN          3                      // pairs of units finished by one epilogue thread
#define KLM_TC_MAX_PAIRS    10                     // N <= 80 columns per CTA
#define KLM_TC_XCHG_BYTES   (2 * (KLM_TC_MAX_PAIRS / 2) * 8 * 64 * 4)
// shared memory before the B image: ring of S chunks (the hand-over buffer aliases it when it fits), barriers
#define KLM_TC_XCHG_OFS     0
#define KLM_TC_FIXED_BYTES(S) ((((S) * KLM_TC_CHUNK_BYTES > KLM_TC_XCHG_BYTES) ? (S) * KLM_TC_CHUNK_BYTES : KLM_TC_XCHG_BYTES) + 128)
#define KLM_TC_TMEM_COLS    256
// single-CTA kernel (klm_seq_tc.cu): up to 12 pairs per CTA; the region in front of the B image holds the ring of
// A-operand chunks and, aliased on it (the ring is idle while the epilogue runs), the hi<->lo hand-over buffer and
// the staging area of the hidden-vector publication ([groups of 8 units][128 rows][8] halves)
#define KLM_TC1_MAX_PAIRS   12
#define KLM_TC1_XCHG_BYTES  (2 * (KLM_TC1_MAX_PAIRS / 2) * 8 * 64 * 4)
#define KLM_TC1_PUB_OFS     KLM_TC1_XCHG_BYTES
#define KLM_TC1_PUB_BYTES   (KLM_TC1_MAX_PAIRS / 4 * 2048)
#define KLM_TC1_ALIAS_BYTES (KLM_TC1_PUB_OFS + KLM_TC1_PUB_BYTES)
// streamed-weights form of the single-CTA kernel (models whose B images do not fit shared memory, e.g. width 1024):
// up to 24 pairs per CTA finished in two epilogue rounds of <= 12; every ring slot holds the A chunk and, behind it,
// the matching k-rows of the CTA's B image; hand-over buffer and publication staging sit in FRONT of the ring (not
// aliased), so that the B parts of the next tick's first chunks can be fetched while the epilogue still runs
// (KlmTcRole::ring_ofs = KLM_TC1_PUB_OFS + 2 KB per group of 8 units of a REC slice)
#define KLM_TCS_MAX_PAIRS   24
#define KLM_TW_MAX_STREAMS  4096                   // wide form (klm_seq_tw): streams of one launch, queued on its 128 rows
#define KLM_TC_TRACE_TICKS  16
// fp16 has 5 exponent bits: the lo part of a hi/lo split (|lo| <= 2^-11 |x|) falls into the subnormal range (below
// 6.1e-5) for |x| < 0.125, and with Keras-default initialisation hidden values are ~1e-3 and weights ~0.05 -- the split
// then keeps 14 instead of 22 bits (states 1.8e-4 off after 20 000 positions, tests/test_gpu_long.py).  Operands are
// therefore scaled by powers of two before the split (exact), hidden values by KLM_TC_HSCALE (|h| < 1), every weight
// matrix by the largest power of two that keeps it below KLM_TC_WMAX, and the fp32 accumulator is scaled back.
#define KLM_TC_HSCALE       4096.f
#define KLM_TC_WMAX         16384.f

// dataflow kernel: ring depths and the layout of KlmTcParams::sync (32-bit words)
#define KLM_DF_R        4                      // versions of a layer's hidden vector kept (hx [D][R][W/8][128][8] halves)
#define KLM_DF_NZ       4                      // slots of Z per layer ([D][NZ][2 parts][W][64] float4)
#define KLM_DF_NL       16                     // slots of the logits ([NL][2 parts][Vp/4][64] float4)
#define KLM_DF_MAXG     4                      // step mode: hypothesis groups of 64 per launch
#define KLM_DF_HFLAG    0                      // [KLM_MAX_DEPTH][64] REC CTA j of layer l: version of its slice that is published
#define KLM_DF_HFLAG0   (1 * KLM_MAX_DEPTH * 64)   // [..][64] step mode: its initial hidden vectors are published
#define KLM_DF_ZPROG    (2 * KLM_MAX_DEPTH * 64)   // [..][64] IN CTA j of layer l: ticks whose slice of Z is complete
#define KLM_DF_ZRD      (3 * KLM_MAX_DEPTH * 64)   // [..][64] REC CTA j of layer l: ticks whose Z it has consumed
#define KLM_DF_HRD      (4 * KLM_MAX_DEPTH * 64)   // [..][64] reader j (IN l+1 / PROJ CTA) of layer l's hidden vector: ticks read
#define KLM_DF_PPROG    (5 * KLM_MAX_DEPTH * 64)   // [64] PROJ CTA j: ticks whose logits are complete
#define KLM_DF_LPROG    (KLM_DF_PPROG + 64)        // [64] LSE warp j: ticks consumed
#define KLM_DF_WORDS    (KLM_DF_LPROG + 64)
// (one word per CTA, never a shared counter: CTAs of one stage may be several ticks apart, a sum would not say that the
//  slowest one is through)

#define KLM_TC_REC  0
#define KLM_TC_IN   1
#define KLM_TC_PROJ 2
#define KLM_TC_LSE  3

struct KlmTcRole {
    int kind;          // KLM_TC_*
    int layer;         // LSTM layer (REC, IN)
    int src_layer;     // whose hidden vector is the A operand
    int sigma;         // wavefront stage: works on position tick - sigma
    int first_pair;    // first pair of units (REC/IN), first group of 8 vocabulary rows (PROJ), LSE index (LSE)
    int npairs;        // pairs owned (N = 8 * npairs columns); number of LSE CTAs (LSE)
    long long wofs;    // byte offset of the CTA's B image in wimg
    int npm;           // CTA-pair kernel: pairs the B image is padded to (the larger of the two CTAs of the cluster)
    // single-CTA kernel: geometry of this CTA's operand ring and how it publishes its hidden-vector slice
    int kc;            // k per ring chunk (chunk = kc * 256 bytes)
    int stages;        // ring slots
    int ring_bytes;    // bytes in front of the barriers (>= stages * chunk and >= KLM_TC1_ALIAS_BYTES)
    int pub;           // REC: 0 = scattered 4-byte stores, 1 = staged in shared memory + coalesced 16-byte stores,
                       //      2 = staged + one bulk async store (1 and 2 need slices aligned to 8 units)
    int ring_ofs;      // streamed-weights form: byte offset of the ring (behind the hand-over buffer and, REC only, the
                       // publication staging of this CTA's slice)
    float scale;       // accumulator -> pre-activation: 1 / (KLM_TC_HSCALE * weight scale of this CTA's matrix), a power of two
    // dataflow kernel (klm_seq_df.cu): who this CTA waits for and who waits for it
    int rot;           // index of the role in plan order (NOT the block it runs in: roles are placed on blocks by measured SM speed);
                       // decides the order in which the CTA walks the operand chunks, so results do not depend on the placement
    int idx;           // index of this CTA within its stage (REC: index of its hand-over flag)
    int n_src;         // REC CTAs of the source layer (flags watched by the producer warp)
    int n_readers;     // REC: CTAs of other stages that read this layer's hidden vector (IN l+1 or PROJ)
    int n_zprod;       // REC l > 0: IN CTAs of this layer (producers of Z)
    int n_zcons;       // IN: REC CTAs of this layer (consumers of Z)
    int alias_slot;    // REC: ring slot that holds the hi/lo hand-over buffer and the publication staging while the ring is idle:
                       //      the slot of the chunk with this CTA's own slice, whose refill is gated by the CTA's own flag
    int guard;         // 1: the aliased buffers do not fit that slot (or step mode): the producer waits for the epilogue instead
    unsigned long long ch_mask[16];   // per operand chunk (natural order): flags of the source-layer CTAs whose slices lie in it
};

// Per-hypothesis arguments of a beam step, carried INSIDE the kernel parameters (constant bank: there when the kernel
// starts -- no copy in front of the launch, no first-touch latency on the step's critical path)
#define KLM_STEP_MAX (64 * 4)                       // = KLM_TC_STREAMS * KLM_DF_MAXG
struct KlmStepArgs {
    int chars[KLM_STEP_MAX], targets[KLM_STEP_MAX], slots_in[KLM_STEP_MAX], slots_out[KLM_STEP_MAX];
    int ctx[KLM_STEP_MAX * KLM_MAX_CTX];
};

// One beam step posted to a RESIDENT instance of klm_seq_df (step server: the kernel stays on the device for all steps of a
// lattice edge / page; weights, TMEM and instruction caches stay warm, a step costs no launch).  The host fills the block in
// mapped page-locked memory and writes `seq` last; the first LSE CTA copies it to device memory and rings the device-side
// doorbell (`seq` of the device copy) for the other CTAs.
struct KlmStepCmd {
    unsigned seq;                 // 1, 2, ...: command number (written last)
    int n_hyp;                    // hypotheses of the step (1 .. KLM_STEP_MAX); < 0: leave the kernel
    int groups;                   // ceil(n_hyp / 64)
    unsigned epoch, tick_base;    // as KlmTcParams::epoch / tick_base of a launch
    int pad_[3];
    KlmStepArgs args;
};

struct KlmTcParams {
    int W, D, V, Vp, Nc;
    int B, T_max, stages;
    int ctx_rows[KLM_MAX_CTX];
    const float4 *P;
    const float4 *Q[KLM_MAX_CTX];
    const float4 *bias[KLM_MAX_DEPTH];
    const unsigned char *wimg;
    const KlmTcRole *roles;
    __half *hx;                   // [D][2][W/8][128][8]
    int kc;                       // CTA-pair kernel: k per ring chunk (64 or 128)
    int ring_bytes;               // CTA-pair kernel: bytes of the ring region (>= the hand-over buffer aliased on it)
    int ingest_bulk;              // CTA-pair kernel only (the single-CTA kernel always uses bulk copies)
    int consumer_fence;           // single-CTA kernel: fence.proxy.async between the grid barrier and the bulk reads
    float *Z;                     // pre-activations of the IN stages: [D][slots][W][64] float4 (i,f,c,o)
    float *logits;                // [2][Vp/4][64] float4
    float *state;
    const int *slots_in, *slots_out;
    const long long *offsets;
    const int *chars, *targets, *n_real, *ctx;
    float *out;
    int full_rows;
    unsigned int *barrier;
    int *err;                     // device-side watchdog: non-zero = a wait timed out (code says which)
    int *err_host;                // dataflow kernel: where a set watchdog word is mirrored at the end of the launch (mapped host memory), or NULL
    unsigned long long *trace;    // optional [grid][KLM_TC_TRACE_TICKS][8] SM-clock timeline (klm_tc_trace)
    int trace_t0;
    int poll_mode;                // grid barrier: 0 = ld.acquire per poll, 1 = relaxed polls + one acquire fence, 2 = relaxed, no fence (experiment)
    int debug_skip;               // timing experiments only (results are wrong): bit 0 = no MMAs, bit 1 = no operand copies
    int trace_mode;               // 0: phase events; 1: arrival of chunk c in event slot c; 2: issue of the copy of chunk c;
                                  // 3: producer steps; 4: epilogue phases of one thread (accumulators complete, per round: drained,
                                  //    hand-over read, finished; arrive)
    int skew;                     // ticks between the REC stages of consecutive layers: 2 = every CTA signals the grid barrier at the
                                  //   end of its tick; 3 = IN / PROJ CTAs signal as soon as their MMAs have read the operand, their
                                  //   epilogue runs into the next tick and Z / the logits are consumed two ticks after the GEMM
                                  //   (stages: REC l at skew*l, IN l at skew*l - (skew-1), PROJ at skew*D - (skew-1), LSE at skew*D)
    int zmask, zlag;              // Z / logits slots: written at slot (tick+1) & zmask, read at (tick+1-zlag) & zmask
                                  //   (skew 2: zmask 1, zlag 1; skew 3: zmask 3, zlag 2)
    int zpub_early;               // 1: IN / PROJ stage their slice and bulk-store it also with the early signal (KLM_TC_ZPUB=2)
    int pub_release;              // 1: a REC CTA signals with red.release.gpu after its bulk store (default 0: the store has been
                                  //    waited for, a relaxed add follows it; KLM_TC_PUB_RELEASE=1 restores the release)
    int stream_b;                 // 1: streamed-weights form (B chunks travel through the ring with the A chunks)
    // ---- dataflow kernel (klm_seq_df.cu): no grid barrier, every hand-over has its own flag / counter ----
    unsigned int *sync;           // flags and counters (layout: KLM_DF_* below); never reset between launches:
    unsigned epoch;               //   flag values of this launch start above `epoch`, counters above n * `tick_base`
    unsigned tick_base;           //   ticks of all earlier launches
    __half *hx0;                  // step mode: initial hidden vectors of all hypothesis groups [D][KLM_DF_MAXG][W/8][128][8]
    int step_groups;              // 0: sequence mode (position = tick); G > 0: step mode -- ONE position for up to 64 * G
                                  //    hypotheses, tick g advances hypotheses 64 g .. 64 g + 63 (state in / out through the pool)
    int n_hyp;                    // step mode: number of hypotheses
    int step_inline;              // step mode: 1 = chars / targets / slots / ctx are taken from `step` below, not from the pointers
    int n_proj, n_lse_warps;      // PROJ CTAs; LSE warps at work
    int exp;                      // experiment switches (klm_set_option "tc_exp"): 1 = back-off between flag polls, 2 = a REC CTA fetches
                                  //   nothing of the next tick before its own publication is through, 4 = IN / PROJ publish at once
    unsigned spin_limit;          // device-side watchdog: polls of one wait before it gives up (klm_set_option "tc_spin_limit")
    int fault;                    // test hook (klm_set_option "tc_fault"): CTA 0 withholds its hand-over signal of tick 2
    int wide;                     // 1: klm_seq_tw -- 128 streams per launch (exchange buffers and LSE rows of 128 streams)
    int server;                   // klm_seq_df, step mode: 1 = resident instance, steps arrive as KlmStepCmd
    const KlmStepCmd *cmd_host;   //   mapped page-locked host memory (read by the first LSE CTA only)
    KlmStepCmd *cmd_dev;          //   device copy; chars / targets / slots_in / slots_out / ctx point into its args
    unsigned *srv_done;           //   mapped host memory: word i = number of the last step LSE CTA i has finished
    const int *lane_ptr;          // klm_seq_tw: row r of the accumulator rates the streams lane_streams[lane_ptr[r] .. lane_ptr[r+1]) one
    const int *lane_streams;      //   after the other, a gap tick between two texts (129 / n ints)
    KlmStepArgs step;             // (last: only the launches of a beam step fill it)
};
