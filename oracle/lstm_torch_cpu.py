"""Batched torch-CPU restatement of the rating arithmetic: the STRONGER of the two CPU baselines.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``): used by ``tests/`` (checked against
``lstm_numpy``) and by ``bench.py``'s CPU-baseline leg, never by the product.

SURVEY.md 8(d) asks for a second CPU line next to the oracle port with the reference's call
structure (batch 1, one model call per window: rating.py:516), "so the speed-up is not inflated
by a slow baseline": here all documents advance together, one ``addmm`` of ``[B, D_l + W] x
[D_l + W, 4W]`` per layer and tick on all host cores (MKL / oneDNN behind torch), fused
projection + log-softmax + target pick.  Same network as ``lstm_numpy`` (rating.py:103-168):
gate order i|f|c|o, exact sigmoid, tied softmax without bias.
"""
import numpy as np
import torch


class TorchCpuModel(object):
    """Weights of a ``lstm_numpy.ModelWeights`` as torch CPU tensors (fp32: the timed baseline; fp64: the truth the
    long parity runs compare with -- bench.py's in-line parity check, tests/test_gpu_long.py); K and R stacked per layer."""

    def __init__(self, w, dtype=np.float32):
        self.dtype = torch.float64 if np.dtype(dtype) == np.float64 else torch.float32
        f = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=dtype))
        self.E = f(w.E)
        self.Et = self.E.t().contiguous()
        self.ctx = [f(c) for c in w.ctx]
        self.KR = [f(np.concatenate([w.K[l], w.R[l]], axis=0)) for l in range(w.depth)]   # [D_l + W, 4W]
        self.b = [f(v) for v in w.b]
        self.width, self.depth, self.voc_size = w.width, w.depth, w.voc_size

    def zero_states(self, n):
        return [torch.zeros(n, self.width, dtype=self.dtype) for _ in range(2 * self.depth)]

    @torch.no_grad()
    def rate(self, x, zs, targets, states=None):
        """Teacher-forced run.  x, targets: int64 [B, T]; zs: list of int64 [B] (one row per stream).

        Returns (log p of the targets [B, T] (model dtype), states list h1,c1,h2,c2,.. of [B, W])."""
        B, T = x.shape
        W = self.width
        x = torch.as_tensor(x, dtype=torch.int64)
        targets = torch.as_tensor(targets, dtype=torch.int64)
        states = states if states is not None else self.zero_states(B)
        h = [states[2 * l] for l in range(self.depth)]
        c = [states[2 * l + 1] for l in range(self.depth)]
        # the context part of layer 1's input does not change over the ticks
        cpart = [tab[torch.as_tensor(z, dtype=torch.int64)] for tab, z in zip(self.ctx, zs)]
        out = torch.empty(B, T, dtype=self.dtype)
        for t in range(T):
            inp = torch.cat([self.E[x[:, t]]] + cpart, dim=1)
            for l in range(self.depth):
                a = torch.addmm(self.b[l], torch.cat([inp, h[l]], dim=1), self.KR[l])
                i, f, g, o = a[:, :W], a[:, W:2 * W], a[:, 2 * W:3 * W], a[:, 3 * W:]
                c[l] = torch.sigmoid(f) * c[l] + torch.sigmoid(i) * torch.tanh(g)
                h[l] = torch.sigmoid(o) * torch.tanh(c[l])
                inp = h[l]
            logits = inp @ self.Et
            out[:, t] = logits.gather(1, targets[:, t:t + 1]).squeeze(1) - torch.logsumexp(logits, dim=1)
        st = []
        for l in range(self.depth):
            st.extend([h[l], c[l]])
        return out.numpy(), st
