"""numpy restatement of the network that ``Rater.configure`` builds.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

Follows /root/reference/ocrd_keraslm/lib/rating.py:
  * embeddings + concat               :103-115, :125
  * stacked LSTM, gate order i|f|c|o, exact sigmoid recurrent activation
                                      :126-145 (``recurrent_activation='sigmoid'`` :133-134)
  * tied softmax  p = softmax(h . E^T), no bias, no projection      :155-168
  * state list order [h1, c1, h2, c2, ...]                          :137-142, :622-629
The per-op arithmetic itself lives in keras==2.3.* (``LSTMCell.call``,
implementation 2: one fused ``x.K + h.R + b`` then split in four) on
tensorflow 1.15 -- third-party, not under /root/reference; restated from the
published algorithm.
"""
import numpy as np


class ModelWeights(object):
    """Plain container: E [V,W]; ctx: list of [rows_k, d_k]; K[l] [D_l,4W]; R[l] [W,4W]; b[l] [4W]."""

    def __init__(self, E, ctx, K, R, b):
        self.E = np.ascontiguousarray(E)
        self.ctx = [np.ascontiguousarray(c) for c in ctx]
        self.K = [np.ascontiguousarray(k) for k in K]
        self.R = [np.ascontiguousarray(r) for r in R]
        self.b = [np.ascontiguousarray(v) for v in b]
        self.width = self.E.shape[1]
        self.voc_size = self.E.shape[0]
        self.depth = len(self.K)
        assert self.K[0].shape[0] == self.width + sum(c.shape[1] for c in self.ctx)
        for l in range(self.depth):
            assert self.K[l].shape[1] == 4 * self.width
            assert self.R[l].shape == (self.width, 4 * self.width)
            assert self.b[l].shape == (4 * self.width,)

    def astype(self, dtype):
        return ModelWeights(self.E.astype(dtype), [c.astype(dtype) for c in self.ctx],
                            [k.astype(dtype) for k in self.K], [r.astype(dtype) for r in self.R],
                            [v.astype(dtype) for v in self.b])

    @property
    def ncontexts(self):
        return len(self.ctx)

    def flops_per_char(self):
        """F = sum_l 2*4W*(D_l+W) + 2*V*W."""
        W = self.width
        return sum(2 * 4 * W * (k.shape[0] + W) for k in self.K) + 2 * self.voc_size * W


def _sigmoid(a):
    # exact logistic, numerically stable on both sides
    out = np.empty_like(a)
    pos = a >= 0
    out[pos] = 1.0 / (1.0 + np.exp(-a[pos]))
    e = np.exp(a[~pos])
    out[~pos] = e / (1.0 + e)
    return out


def embed(w, x, zs):
    """x: int [n]; zs: list of int [n] -> u [n, D1]   (rating.py:103-115,125)."""
    parts = [w.E[x]]
    for tab, z in zip(w.ctx, zs):
        parts.append(tab[z])
    return np.concatenate(parts, axis=-1)


def cell(w, l, inp, h, c):
    """One LSTM layer timestep for n rows (Keras LSTMCell, implementation 2)."""
    W = w.width
    a = inp @ w.K[l] + h @ w.R[l] + w.b[l]
    i = _sigmoid(a[:, 0:W])
    f = _sigmoid(a[:, W:2 * W])
    g = np.tanh(a[:, 2 * W:3 * W])
    o = _sigmoid(a[:, 3 * W:4 * W])
    c_new = f * c + i * g
    h_new = o * np.tanh(c_new)
    return h_new, c_new


def softmax_rows(logits):
    m = logits.max(axis=-1, keepdims=True)
    e = np.exp(logits - m)
    return e / e.sum(axis=-1, keepdims=True)


def step(w, x, zs, states):
    """One timestep for n rows.

    x int[n]; zs list of int[n]; states list of 2*depth arrays [n,W] (h1,c1,h2,c2,..).
    Returns (probs [n,V], new_states list).
    """
    inp = embed(w, x, zs)
    new_states = []
    for l in range(w.depth):
        h, c = cell(w, l, inp, states[2 * l], states[2 * l + 1])
        new_states.extend([h, c])
        inp = h
    probs = softmax_rows(inp @ w.E.T)
    return probs, new_states


def zero_states(w, n, dtype=None):
    dtype = dtype or w.E.dtype
    return [np.zeros((n, w.width), dtype=dtype) for _ in range(2 * w.depth)]


def run_sequence(w, x, zs, states, want='probs', targets=None):
    """Teacher-forced run over T steps for n parallel rows.

    x int[n,T]; zs list of int[n,T]; states as in ``step`` (updated copy returned).
    want='probs' -> [n,T,V]; want='target' -> p of targets[n,T] only, [n,T].
    """
    n, T = x.shape
    dtype = w.E.dtype
    if want == 'probs':
        out = np.empty((n, T, w.voc_size), dtype=dtype)
    else:
        out = np.empty((n, T), dtype=dtype)
    rows = np.arange(n)
    for t in range(T):
        probs, states = step(w, x[:, t], [z[:, t] for z in zs], states)
        if want == 'probs':
            out[:, t, :] = probs
        else:
            out[:, t] = probs[rows, targets[:, t]]
    return out, states
