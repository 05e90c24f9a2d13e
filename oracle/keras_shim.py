"""A stand-in for the ``keras.Model`` object that the reference's ``Rater`` holds in
``self.model``, backed by ``oracle.lstm_numpy``.

TEST INFRASTRUCTURE ONLY.  Used by ``tests/golden/make_golden.py`` (build container
only) to execute the reference's *own, unmodified* host logic
(/root/reference/ocrd_keraslm/lib/rating.py, imported in place with ``h5py``
stubbed) so that golden vectors carry reference-exact windowing, padding, beam
pruning and path selection.  Exposes exactly what rating.py touches on
``self.model``:

  predict_generator(gen, steps=, verbose=)   rating.py:516
  predict_on_batch(list)                     rating.py:566, :631
  evaluate_generator(gen, steps=, verbose=)  rating.py:490
  reset_states()                             rating.py:475, :555
  inputs (objects with .name)                rating.py:183
"""
import types
import numpy as np

from . import lstm_numpy as ln


class _Input(object):
    def __init__(self, name):
        self.name = name


class KerasModelShim(object):
    def __init__(self, weights, stateful, incremental):
        self.w = weights
        self.stateful = stateful
        self.incremental = incremental
        names = ['char_input'] + ['context%d_input' % (k + 1) for k in range(weights.ncontexts)]
        if incremental:
            for l in range(weights.depth):
                names += ['initial_h_%d_input' % (l + 1), 'initial_c_%d_input' % (l + 1)]
        self.inputs = [_Input(n + ':0') for n in names]
        self.states = ln.zero_states(weights, 1)
        self.calls = 0

    def reset_states(self):
        self.states = ln.zero_states(self.w, 1)

    def _run_window(self, xs):
        x = np.asarray(xs[0]).astype(np.int64)
        zs = [np.asarray(z).astype(np.int64) for z in xs[1:1 + self.w.ncontexts]]
        out, self.states = ln.run_sequence(self.w, x, zs, self.states, want='probs')
        return out

    def predict_generator(self, generator, steps=None, verbose=0):
        assert self.stateful
        outs = []
        for _ in range(steps):
            xs, _y = next(generator)
            outs.append(self._run_window(xs))
            self.calls += 1
        return np.concatenate(outs, axis=0)

    def evaluate_generator(self, generator, steps=None, verbose=0):
        # Keras: mean over batches of (mean over all positions of categorical crossentropy),
        # batch size 1; y rows that are all-zero (padding) contribute -sum(0*log p) = 0.
        assert self.stateful
        losses, accs = [], []
        for _ in range(steps):
            xs, y = next(generator)
            p = self._run_window(xs)
            p = np.clip(p / p.sum(axis=-1, keepdims=True), 1e-7, 1 - 1e-7)
            ce = -(y * np.log(p)).sum(axis=-1)
            losses.append(ce.mean())
            accs.append((p.argmax(-1) == y.argmax(-1)).mean())
        return float(np.mean(losses)), float(np.mean(accs))

    def predict_on_batch(self, inputs):
        self.calls += 1
        if self.stateful:
            return self._run_window(inputs)
        assert self.incremental
        nc = self.w.ncontexts
        x = np.asarray(inputs[0]).astype(np.int64)[:, 0]
        zs = [np.asarray(z).astype(np.int64)[:, 0] for z in inputs[1:1 + nc]]
        states = [np.asarray(s).astype(self.w.E.dtype) for s in inputs[1 + nc:]]
        probs, new_states = ln.step(self.w, x, zs, states)
        return [probs] + new_states


def import_reference_rating(path='/root/reference'):
    """Import the reference's rating.py in place (never copied), with h5py stubbed."""
    import sys
    import importlib.util
    if 'h5py' not in sys.modules:
        sys.modules['h5py'] = types.ModuleType('h5py')
    spec = importlib.util.spec_from_file_location('_ref_rating', path + '/ocrd_keraslm/lib/rating.py')
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def make_reference_rater(ref_mod, weights, mapping, length, stateful, incremental, logger=None):
    """Build a reference ``Rater`` whose model is the shim (status 2 = trained)."""
    r = ref_mod.Rater(logger=logger)
    r.width, r.depth, r.length = weights.width, weights.depth, length
    r.stateful, r.incremental = stateful, incremental
    if stateful:
        r.batch_size = 1
        r.variable_length = False
        r.first_window = 0
    r.mapping = mapping
    r.voc_size = weights.voc_size
    r.model = KerasModelShim(weights, stateful, incremental)
    r.status = 2
    return r
