/*
 * keraslm_b200.h -- C ABI of libkeraslm_b200.so: the sm_100a device side of the
 * ocrd_keraslm LSTM rating path.
 *
 * The reference (OCR-D/ocrd_keraslm) has no FFI of its own for this path: everything
 * below `ocrd_keraslm.lib.Rater` is Keras 2.3 / TensorFlow 1.15 calls.  Every entry
 * point here therefore cites the reference *call site* it replaces
 * (file:line in /root/reference/ocrd_keraslm/lib/rating.py unless noted).
 *
 * Conventions
 *   - every function returns 0 on success, a negative KLM_E* code otherwise;
 *     klm_last_error() gives the message of the last failure on the calling thread.
 *   - no exceptions, no torch types: plain pointers and sizes.
 *   - `mem` says where the caller's index/score buffers live: KLM_MEM_HOST (the
 *     library stages them through its own device scratch, H2D/D2H on the model's
 *     stream, and returns after the results are in the host buffer) or
 *     KLM_MEM_DEVICE (pointers are device pointers on the model's device; the call
 *     only enqueues work on the model's stream).
 *   - a model handle is not re-entrant (one Rater <-> one device <-> one stream,
 *     like the reference: wrapper/rate.py:65 `max_workers = 1`).
 *   - LSTM states live in a device-resident pool of `slots`; one slot holds
 *     [depth][2][width] float32 = h1,c1,h2,c2,... (order of rating.py:137-142,622-629).
 */
#ifndef KERASLM_B200_H
#define KERASLM_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KLM_MAX_CTX    4
#define KLM_MAX_DEPTH  8

#define KLM_OK          0
#define KLM_EINVAL     -1   /* bad argument (shape, index out of range, NULL) */
#define KLM_ECUDA      -2   /* CUDA runtime failure, message has the CUDA error string */
#define KLM_ENOMEM     -3
#define KLM_ESTATE     -4   /* slot out of range / pool not allocated */

#define KLM_MEM_HOST    0
#define KLM_MEM_DEVICE  1

typedef struct klm_model klm_model;

/* Shape of the network that Rater.configure builds (rating.py:61-179). */
typedef struct klm_config {
    int32_t width;                  /* LSTM units per layer = char embedding dim (:103-106) */
    int32_t depth;                  /* number of LSTM layers (:126) */
    int32_t voc_size;               /* rows of the char embedding = output classes (:104,:163) */
    int32_t n_ctx;                  /* number of context variables (reference: 1, :110-114) */
    int32_t ctx_rows[KLM_MAX_CTX];  /* rows of each context embedding (reference: 200) */
    int32_t ctx_dim[KLM_MAX_CTX];   /* columns of each context embedding (reference: 10) */
} klm_config;

/*
 * Replaces Rater.configure + Rater.load_weights (rating.py:61-179, :966-974): uploads the
 * float32 weights (host pointers, Keras layouts) and prepares the device-resident
 * forms.  `weights` holds 1 + n_ctx + 3*depth pointers in this order:
 *   E [voc_size, width], C_1 [ctx_rows, ctx_dim] ... C_nctx,
 *   then per layer l: kernel [D_l, 4*width], recurrent_kernel [width, 4*width], bias [4*width]
 *   (gate order i|f|c|o; D_1 = width + sum ctx_dim, D_l>1 = width).
 */
int klm_create(const klm_config *cfg, const float *const *weights, int device, klm_model **out);
int klm_destroy(klm_model *m);

/* Message of the last failure on this thread ("" if none). */
const char *klm_last_error(void);

/* sizeof(klm_config) (what = 0) / sizeof(klm_info) (what = 1) as this library was compiled: lets a binding check its
 * own struct declarations before it passes them in. */
int klm_abi_sizeof(int what);

/* Run subsequent work on this CUDA stream (a cudaStream_t; NULL = the library's own stream). */
int klm_set_stream(klm_model *m, void *cuda_stream);
/* Block until everything enqueued on the model's stream has finished; reports a device-side watchdog error
 * (a hand-over wait inside a persistent kernel that timed out) of any launch so far -- the place where
 * KLM_MEM_DEVICE callers, whose klm_rate / klm_step only enqueue, learn about it. */
int klm_sync(klm_model *m);
/* Run-time switches of the library (diagnostics and timing experiments; the defaults are the product configuration).
 * Initial values come from the environment variables of the same name in upper case with the prefix KLM_
 * (e.g. KLM_TC_SKEW), read once in klm_create.  Names: tc_skew, tc_trace_mode, tc_debug_skip, tc_poll, tc_zpub,
 * tc_pub_release, tc_spin_limit, tc_fault, tc_df, tc_wide, tc_exp, tc_consumer_fence, ll_debug, step_mapped,
 * step_server, step_timing. */
int klm_set_option(klm_model *m, const char *name, int64_t value);

/*
 * State pool.  Replaces the Keras backend variables of the stateful LSTM layers
 * (rating.py:90-92,127-145) and the numpy state lists that Rater.predict passes
 * around (:622-638).  New slots are zero (= model.reset_states(), :475,:555).
 */
int klm_state_pool_resize(klm_model *m, int32_t n_slots);
int klm_state_reset(klm_model *m, const int32_t *slots, int32_t n);                      /* host ints */
int klm_state_copy(klm_model *m, const int32_t *src, const int32_t *dst, int32_t n);     /* host ints */
int klm_state_to_host(klm_model *m, const int32_t *slots, int32_t n, float *out);        /* [n][2*depth][width] */
int klm_state_from_host(klm_model *m, const int32_t *slots, int32_t n, const float *in);
/* L2 distances |state_a[k] - state_b[k]| for the first `n_arrays` arrays of the h1,c1,h2,..
 * list (rating.py:887-916 compares range(depth) of them); out [n][n_arrays] host floats. */
int klm_state_l2_distance(klm_model *m, const int32_t *a, const int32_t *b, int32_t n,
                          int32_t n_arrays, float *out);
/* klm_state_l2_distance that allocates NOTHING (no cudaMalloc / cudaFree, no allocating library
 * call): for callers that cannot allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  It has no counterpart in the reference.
 * Same arguments, same kernel, same results; the scratch (the
 * device copies of the two slot lists and the distances) lives in ONE static device array of
 * KLM_NOALLOC_SCRATCH_BYTES that every GPU which loads the library holds until exit.
 * Limit: 256-byte-rounded 8*n + 4*n*n_arrays <= KLM_NOALLOC_SCRATCH_BYTES (e.g. 2048 pairs x 6
 * arrays); beyond it the call returns KLM_EINVAL and the message names the size.
 * All calls share the array (and it keeps the last call's contents; the part a call reads back
 * is cleared first): calls that overlap on one GPU -- two models, two host threads -- race. */
#define KLM_NOALLOC_SCRATCH_BYTES 65536
int klm_state_l2_distance_noalloc(klm_model *m, const int32_t *a, const int32_t *b, int32_t n,
                                  int32_t n_arrays, float *out);

/* Host-only helper of the sharded form (the reference has no multi-GPU code): longest-processing-time-first
 * assignment of n independent units (documents, pages, lines) to `world` ranks.  Units are taken by decreasing length
 * (ties: lower index first) and given to the least loaded rank (ties: lower rank); a unit costs its length + 1.
 * rank_of[i] receives the rank of unit i.  Deterministic, so every rank computes the same partition without
 * communicating (ocrd_keraslm_b200/shard.py: partition).  No device work, no model handle. */
int klm_lpt_partition(const int64_t *lengths, int64_t n, int32_t world, int32_t *rank_of);

/*
 * Teacher-forced rating of n_streams independent character streams in ONE launch.
 * Replaces model.predict_generator over the windows of Rater._gen_data (rating.py:516,
 * :1005-1102) plus the target pick of Rater.rate (:520-528), and the per-character
 * predict_on_batch loop of Rater.rate2 (:555-575).
 *
 *   slots    [n]      host ints: state slot of each stream; state is read at the start and
 *                     written back at the end (stateful carry, no reset: :493-529)
 *   offsets  [n+1]    stream s owns positions offsets[s] .. offsets[s+1]-1 of chars/targets/out
 *   chars    [total]  input character index per position (0 = unmapped / padding, :1133-1135)
 *   targets  [total]  index whose score is wanted at that position (next character, :522-526)
 *   n_real   [n]      positions >= n_real[s] are window padding: the context index is 0 there
 *                     (:1141-1143); NULL = all positions real
 *   ctx      [n][n_ctx] context indices per stream (time-constant, :1142)
 *   out_logp [total]  natural-log probability of targets[] at each position (float32)
 * offsets is int64, the others int32.  chars, targets and out_logp live where `mem` says; slots, offsets, n_real
 * and ctx are ALWAYS host pointers (they are read on the host to plan the launches).
 */
int klm_rate(klm_model *m, int32_t n_streams, const int32_t *slots, const int64_t *offsets,
             const int32_t *chars, const int32_t *targets, const int32_t *n_real,
             const int32_t *ctx, float *out_logp, int mem);

/*
 * One timestep for n hypotheses with explicit state passing.  Replaces
 * model.predict_on_batch in Rater.predict (rating.py:631) incl. the vstack/slicing
 * of states around it (:625-638).
 *
 *   chars     [n]        input character index per hypothesis
 *   ctx       [n][n_ctx] context indices
 *   slots_in  [n]        host ints; -1 = zero state (None in the reference, :625-627)
 *   slots_out [n]        host ints; distinct; where the new state goes (may equal slots_in[i])
 *   targets   [n] or NULL
 *   out       targets != NULL: [n] natural-log probability of targets[i]
 *             targets == NULL: [n][voc_size] full softmax rows (what predict returns, :632)
 * chars, targets and out live where `mem` says; ctx, slots_in and slots_out are ALWAYS host pointers.
 */
int klm_step(klm_model *m, int32_t n, const int32_t *chars, const int32_t *ctx,
             const int32_t *slots_in, const int32_t *slots_out, const int32_t *targets,
             float *out, int mem);

/*
 * klm_step whose result is reduced on the device: instead of the n x voc_size softmax rows only the k most probable
 * entries of every row come back, ascending by probability -- what Rater.generate keeps of them
 * (rating.py:690-704: `argsort(pred)[-10:]`, then a threshold).  out_index / out_prob: [n][k] host arrays.
 */
int klm_step_topk(klm_model *m, int32_t n, const int32_t *chars, const int32_t *ctx, const int32_t *slots_in,
                  const int32_t *slots_out, int32_t k, int32_t *out_index, float *out_prob);

/*
 * One lattice edge of the beam search.  Replaces the per-edge loop of Rater.rate_best (rating.py:789-852):
 * candidates = incoming hypotheses x alternatives, character-synchronous batched steps (klm_step with
 * targets), worst-first pops, projected-cost ordering with bisect.insort_left semantics, the +15 bit
 * cardinality break (:816), the +2.5 bit threshold (:847) and the truncation (:851).  Not covered (the caller
 * keeps the reference control flow for these): beam_clustering_dist != 0, an out-node that already has a beam.
 *
 *   hyp_slot/hyp_last_char/hyp_cum_cost [n_hyp]  state slot (-1 = zeros), index of the last character and
 *                                                cumulative cost of every hypothesis of the incoming beam
 *   alt_offsets [n_alt+1], alt_chars             character indices of the alternatives, alt_conf [n_alt]
 *   ctx [n_ctx]                                  context indices; lm_weight, batch_size as in the reference
 *   scratch_slots [n_scratch]                    free state slots the search may use (>= n_hyp*n_alt + batch_size)
 *   out_* [max_out]                              the first max_out finished candidates in beam order: hypothesis
 *                                                index, alternative index, cumulative cost, state slot (one of
 *                                                scratch_slots, or -2 = the incoming hypothesis' own state)
 * All pointers are host pointers; the call returns when the edge is finished.
 */
int klm_beam_edge(klm_model *m, int32_t n_hyp, const int32_t *hyp_slot, const int32_t *hyp_last_char,
                  const double *hyp_cum_cost, int32_t n_alt, const int32_t *alt_offsets, const int32_t *alt_chars,
                  const double *alt_conf, const int32_t *ctx, double lm_weight, int32_t batch_size,
                  int32_t n_scratch, const int32_t *scratch_slots, int32_t max_out, int32_t *out_hyp,
                  int32_t *out_alt, double *out_cum_cost, int32_t *out_slot, int32_t *n_out, int64_t *n_steps,
                  int64_t *n_hyp_steps);

/*
 * A whole LINEAR lattice in one call -- the chain of edges wrapper/rate.py:343-424 builds from a PAGE: klm_beam_edge's
 * search edge after edge, the first `beam_width` finished candidates of an edge being the incoming hypotheses of the next
 * (rating.py:763-852 for a graph without branches and without history clustering).  Replaces one Python round trip per
 * edge; the caller builds its back-pointer tree from the table it gets.
 *
 *   edge_alt_first [n_edges+1]                   alternatives edge_alt_first[e] .. edge_alt_first[e+1]-1 belong to edge e
 *   alt_offsets [n_alts+1], alt_chars, alt_conf  character indices and confidences of all alternatives of all edges
 *   out_first [n_edges+1]                        the kept candidates of edge e are rows out_first[e] .. out_first[e+1]-1 of
 *   out_hyp / out_alt / out_cum_cost [max_out]   (row of the previous edge -- for edge 0: incoming hypothesis --,
 *                                                alternative index within the edge, cumulative cost), in beam order
 *   final_slot [beam_width]                      state of each candidate of the last edge: one of scratch_slots (now the
 *                                                caller's), or -2 - i = the untouched state of incoming hypothesis i
 *   scratch_slots [n_scratch]                    free state slots: >= beam_width * (most alternatives of an edge + 1) + batch_size
 * All pointers are host pointers; the call returns when the lattice is finished.
 */
int klm_beam_lattice(klm_model *m, int32_t n_hyp, const int32_t *hyp_slot, const int32_t *hyp_last_char,
                     const double *hyp_cum_cost, int32_t n_edges, const int32_t *edge_alt_first,
                     const int32_t *alt_offsets, const int32_t *alt_chars, const double *alt_conf, const int32_t *ctx,
                     double lm_weight, int32_t batch_size, int32_t beam_width, int32_t n_scratch,
                     const int32_t *scratch_slots, int32_t max_out, int32_t *out_first, int32_t *out_hyp,
                     int32_t *out_alt, double *out_cum_cost, int32_t *final_slot, int64_t *n_steps, int64_t *n_hyp_steps);

/* Introspection for bench.py / tests. */
typedef struct klm_info {
    int32_t sm_count;
    int32_t grid;               /* CTAs of the persistent kernel */
    int32_t weights_in_smem;    /* 1 if every CTA keeps its weight slice in shared memory */
    int32_t smem_bytes;         /* dynamic shared memory per CTA */
    int64_t kernel_launches;    /* kernels of this library launched so far on this model */
    double  last_kernel_ms;     /* device time of the last klm_rate/klm_step main kernel (CUDA events) */
    int32_t max_streams;        /* largest n one launch takes (4096 with tc_wide: queued on 128 rows); larger calls are split */
    int32_t tc_enabled;         /* tcgen05 sequence kernel serving 9..64 streams: 0 none (FMA kernels), 1 weights resident in
                                 * shared memory, 2 CTA-pair variant, 3 weights streamed through the operand ring */
    int32_t tc_grid;            /* CTAs of the tcgen05 kernel */
    int32_t tc_smem_bytes;      /* dynamic shared memory per CTA of the tcgen05 kernel */
    int64_t tc_launches;        /* launches of the tcgen05 kernel so far */
    int64_t cl_launches;        /* launches of the one-cluster kernel (<= 8 streams, small models) so far */
    int32_t cl_cluster;         /* CTAs of that cluster (8 or 16); 0 = the model does not fit one cluster */
    int32_t tc_dataflow;        /* 1: the tcgen05 kernel in use is the dataflow form (a flag per hand-over, no grid barrier;
                                 * also serves klm_step for up to 256 hypotheses per launch), 0: the grid-barrier form */
    int32_t tc_wide;            /* 1: calls of more than 64 streams run in launches of up to 4096 queued on 128 rows (klm_seq_tw, streamed-weights plans) */
    int64_t tc_wide_launches;   /* launches of klm_seq_tw so far (counted in tc_launches as well) */
} klm_info;
int klm_get_info(klm_model *m, klm_info *out);

/* Measured FP32 FMA throughput of the device in TFLOP/s (a ~2 ms micro-benchmark: independent FFMA chains on every SM):
 * the roofline denominator of the small-batch kernels (<= 8 streams), which run on the FP32 pipes, not the tensor cores. */
int klm_measure_fma_peak(int device, double *tflops);

/* Self-test of the tcgen05 / TMEM / bulk-copy building blocks: D[128 x N] = A[128 x K] . B[N x K]^T for
 * fp16 operand images in the kernel's shared-memory layout ([K/8][rows][8] halves); host pointers. */
int klm_selftest_umma(int device, int32_t N, int32_t K, const uint16_t *A_img, const uint16_t *B_img, float *D);

/* Self-test of the CTA-pair MMA (cta_group::2, M = 128 over two CTAs): A_img = the two [K/8][64][8] row halves,
 * B_img = the two [K/8][N/2][8] row halves; D = raw TMEM dump [2 CTAs][128 lanes][N/2 columns]; host pointers. */
int klm_selftest_umma2(int device, int32_t N, int32_t K, const uint16_t *A_img, const uint16_t *B_img, float *D);

/* Debug timeline of the tcgen05 kernel.  out == NULL: arm tracing of 16 ticks from `first_tick` for the
 * following launches; out != NULL: copy [tc_grid][16][8] SM-clock samples (events: 0 tick start, 1 copies
 * issued, 2 first / 3 last operand chunk landed, 4 accumulators complete, 5 TMEM drained, 6 hand-over stored,
 * 7 barrier arrive) and, if roles_out != NULL, [tc_grid][4] ints (kind, layer, pairs, stage). */
int klm_tc_trace(klm_model *m, int32_t first_tick, uint64_t *out, int32_t *roles_out);

#ifdef __cplusplus
}
#endif
#endif /* KERASLM_B200_H */
