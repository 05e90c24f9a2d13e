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
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/keraslm_b200.h"

#define KLM_THREADS      512
#define KLM_WARPS        (KLM_THREADS / 32)
#define KLM_MAX_ITEMS    8      // (unit, 32-stream group) items per warp in the stream-parallel mode
#define KLM_MAX_UNITS_K  4      // units per warp in the K-parallel mode
#define KLM_MODEK_MAXB   8      // K-parallel mode takes up to this many streams

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
};
