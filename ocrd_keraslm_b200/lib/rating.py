"""B200-native ``Rater`` / ``Node``: the drop-in for ``ocrd_keraslm.lib`` on the rating path.

Same public surface and call order as the reference class
(/root/reference/ocrd_keraslm/lib/rating.py: ``Rater`` :22-1238, ``Node`` :1240-1318):

    r = Rater(logger=...)            :35
    r.load_config(filename)          :947   (status must be 0)
    r.stateful / r.incremental = ..  wrapper/rate.py:84-88
    r.configure()                    :61    (status 1)
    r.load_weights(filename)         :966   (status 2)
    r.rate / rate2 / predict / rate_best / next_path / generate / test

Everything the reference delegated to Keras/TensorFlow (embedding gather, stacked LSTM, tied
softmax, state variables) runs in hand-written sm_100a kernels behind the C ABI of
``include/keraslm_b200.h``; there is no other backend and no CPU fallback.  LSTM states live in a
device-resident pool; what ``predict`` hands back are ``DeviceState`` handles that behave like the
reference's lists of ``(1, width)`` arrays when somebody looks inside.
"""
import bisect
import json
import logging
import math
import os
import unicodedata

import numpy as np

from .. import _native
from .. import h5lite


class Node(object):
    """Back-pointer tree node of the beam search (reference :1240-1318).

    Ordering *and equality* are by projected cost -- cumulative cost plus 0.5 bit per character
    still missing from the alternative this node is spelling out (:1300-1318); nodes are
    therefore unhashable, and ``in`` / ``list.remove`` match by cost, which callers rely on.
    """
    __hash__ = None

    def __init__(self, state, value, cost, parent=None, extras=None):
        self.value = value
        self.parent = parent
        self.state = state
        self.cum_cost = parent.cum_cost + cost if parent else cost
        self.length = parent.length + 1 if parent is not None else 1
        self.extras = extras
        self._sequence = None

    def to_sequence(self, stop_at=None):
        """Nodes from the root (or from the nearest ancestor found in `stop_at`) down to here; cached."""
        if not self._sequence:
            seq = []
            take = not stop_at
            cur = self
            while cur:
                if stop_at and cur in stop_at:
                    take = True
                if take:
                    seq.insert(0, cur)
                cur = cur.parent
            self._sequence = seq
        return self._sequence

    def cut_at(self, node):
        """Make the ancestor whose parent is `node` the new root of this history."""
        cur = self
        while cur:
            if cur.parent is node:
                cur.parent = None
                self._sequence = None
                break
            cur = cur.parent

    def pro_cost(self):
        todo = len(self.extras[1].Unicode) - len(self.value) if self.extras else 0
        return self.cum_cost + 0.5 * todo

    def __lt__(self, other):
        return self.pro_cost() < other.pro_cost()

    def __le__(self, other):
        return self.pro_cost() <= other.pro_cost()

    def __eq__(self, other):
        return self.pro_cost() == other.pro_cost()

    def __ne__(self, other):
        return self.pro_cost() != other.pro_cost()

    def __gt__(self, other):
        return self.pro_cost() > other.pro_cost()

    def __ge__(self, other):
        return self.pro_cost() >= other.pro_cost()


class DeviceState(object):
    """Handle of one LSTM state (h1,c1,h2,c2,...) held in the device pool.

    Looks like the reference's per-hypothesis state list (:636-638): ``len()`` is 2*depth and
    item k is a float32 array of shape (1, width), fetched from the device on first access.
    The slot returns to the pool when the handle is garbage collected.
    """
    __slots__ = ('_pool', 'slot', '_host')

    def __init__(self, pool, slot):
        self._pool = pool
        self.slot = slot
        self._host = None

    def __del__(self):
        try:
            self._pool.release(self.slot)
        except Exception:
            pass

    def __len__(self):
        return 2 * self._pool.depth

    def _fetch(self):
        if self._host is None:
            self._host = self._pool.model.state_to_host([self.slot])[0]
        return self._host

    def __getitem__(self, k):
        host = self._fetch()
        if isinstance(k, slice):
            return [host[i:i + 1] for i in range(*k.indices(len(host)))]
        return host[k][None, :]

    def __iter__(self):
        host = self._fetch()
        return iter([host[i:i + 1] for i in range(len(host))])


class _StatePool(object):
    """Free-list over the slots of the native state pool."""

    def __init__(self, model, depth, reserve=1):
        self.model = model
        self.depth = depth
        self.capacity = 0
        self.free = []
        self.reserved = reserve            # slots [0, reserve) belong to the stateful Rater itself
        self._grow(max(64, reserve))

    def _grow(self, n):
        self.model.pool_resize(n)
        lo = max(self.capacity, self.reserved)
        self.free.extend(range(n - 1, lo - 1, -1))
        self.capacity = n

    def acquire(self, n):
        if len(self.free) < n:
            self._grow(max(2 * self.capacity, self.capacity + n))
        if n <= 0:
            return []
        out = self.free[:-n - 1:-1]                # the same slots, in the same order, as n pops
        del self.free[len(self.free) - n:]
        return out

    def release(self, slot):
        self.free.append(slot)

    def release_many(self, slots):
        self.free.extend(slots)


class _Scores(list):
    """The list `rate_many` returns; `flat` (if not None) is the concatenation of its arrays, which are views of it."""
    flat = None


class _DeviceModelView(object):
    """What sits in ``Rater.model``: the handful of Keras-model methods callers may touch."""

    def __init__(self, rater):
        self._rater = rater
        self.inputs = []

    def reset_states(self):
        self._rater.reset_states()


class Rater(object):
    """Rate character sequences with a stacked-LSTM character language model (reference :22-59)."""

    def __init__(self, logger=None, device=None):
        # configuration (overwritten by load_config)
        self.width = 0
        self.depth = 0
        self.length = 0
        self.variable_length = True
        self.first_window = 0.1
        self.char_degradation = 0.01
        self.context_degradation = 0.1
        self.stateful = True
        self.mapping = ({}, {})
        self.batch_size = 128
        self.validation_split = 0.2
        self.smoothing = 0.2
        # runtime
        self.logger = logger or logging.getLogger(__name__)
        self.reset_cb = None
        self.incremental = False
        self.model = None
        self.history = {}
        self.status = 0
        self.voc_size = 0
        # device side
        if device is None:
            device = int(os.environ.get('LOCAL_RANK', '0')) if 'KLM_DEVICE' not in os.environ \
                else int(os.environ['KLM_DEVICE'])
        self.device = device
        self.context_shapes = [(200, 10)]       # reference: one table, year/10 (:110-114)
        self._native = None
        self._pool = None
        self._lut = None
        self._lut_src = None
        self._pending_weights = None
        self._pin = None
        self._worker = None
        self.native_beam = True            # rate_best: run whole lattice edges through klm_beam_edge where possible
        self.native_lattice = True         # ... and whole linear lattices (a PAGE) through klm_beam_lattice
        self.device_steps = 0              # klm_step calls / hypotheses advanced so far (bench.py: beam steps/sec)
        self.hypothesis_steps = 0

    # ------------------------------------------------------------------ lifecycle
    def load_config(self, filename):
        """Read /config of the model file (reference :947-964)."""
        assert self.status == 0
        with h5lite.File(filename, 'r') as f:
            group = f['config']
            self.history = json.loads(_as_str(group['history'][()])) if 'history' in group else {}
            self.width = int(group['width'][()])
            self.depth = int(group['depth'][()])
            self.length = int(group['length'][()])
            self.stateful = bool(group['stateful'][()])
            self.variable_length = bool(group['variable_length'][()])
            self.set_mapping(group['mapping'][()])
            # context tables are not part of /config: take their shapes from the stored weights
            shapes = []
            k = 1
            while 'context%d_embedding' % k in f:
                shapes.append(tuple(int(x) for x in _layer_weights(f, 'context%d_embedding' % k)[0].shape))
                k += 1
            if shapes:
                self.context_shapes = shapes

    def set_mapping(self, mapping_array):
        """Vocabulary from the stored uint32 code point array (index = vocabulary id, 0 = unmapped; :961-964)."""
        arr = np.asarray(mapping_array)
        c_i = dict((chr(int(c)), i) for i, c in enumerate(arr) if c > 0)
        i_c = dict((i, chr(int(c))) for i, c in enumerate(arr) if c > 0)
        self.mapping = (c_i, i_c)
        self.voc_size = len(c_i) + 1

    def configure(self):
        """Fix the mode of operation (reference :61-179).  The network itself is built on the
        device by load_weights / set_weights, since its kernels are specialised by the weights."""
        if self.stateful:
            self.batch_size = 1                 # :91
            self.variable_length = False        # :73-76
            self.first_window = 0
        nctx = sum(d for _, d in self.context_shapes)
        self.logger.info('Compiling %s %s model with depth %d width %d length %d context size %d',
                         'sm_100a', 'stateful' if self.stateful else
                         ('incremental' if self.incremental else 'stateless'),
                         self.depth, self.width, self.length, nctx)
        self.model = _DeviceModelView(self)
        names = ['char_input'] + ['context%d_input' % (k + 1) for k in range(len(self.context_shapes))]
        self.model.inputs = [_Named(n + ':0') for n in names]
        self.status = 1
        if self._pending_weights is not None:
            self.set_weights(self._pending_weights)

    def load_weights(self, filename):
        """Load the layer weights of a Keras-format model file onto the device (reference :966-974)."""
        assert self.status > 0
        with h5lite.File(filename, 'r') as f:
            wd = read_keras_weights(f, self.depth)
        self.set_weights(wd)

    def set_weights(self, wd):
        """Weights as {'char_embedding': E, 'context1_embedding': C, 'lstm_1': (K, R, b), ...}."""
        assert self.status > 0
        E = np.asarray(wd['char_embedding'], dtype=np.float32)
        ctx = []
        k = 1
        while 'context%d_embedding' % k in wd:
            ctx.append(np.asarray(wd['context%d_embedding' % k], dtype=np.float32))
            k += 1
        layers = []
        for l in range(self.depth):
            K, R, b = wd['lstm_%d' % (l + 1)]
            K, R, b = convert_cudnn_weights(np.asarray(K, np.float32), np.asarray(R, np.float32),
                                            np.asarray(b, np.float32))
            layers += [K, R, b]
        W = self.width
        if E.shape[1] != W:
            raise ValueError('char embedding has dim %d but width is %d' % (E.shape[1], W))
        d_in = W + sum(c.shape[1] for c in ctx)
        for l in range(self.depth):
            exp_in = d_in if l == 0 else W
            if layers[3 * l].shape != (exp_in, 4 * W) or layers[3 * l + 1].shape != (W, 4 * W) or \
                    layers[3 * l + 2].shape != (4 * W,):
                raise ValueError('lstm_%d weights have shapes %s, expected (%d|%d, %d)' % (
                    l + 1, [a.shape for a in layers[3 * l:3 * l + 3]], exp_in, W, 4 * W))
        self.context_shapes = [tuple(c.shape) for c in ctx]
        self._host_weights = wd
        self._native = _native.Model(W, self.depth, E.shape[0], self.context_shapes, [E] + ctx + layers,
                                     device=self.device)
        self._pool = _StatePool(self._native, self.depth, reserve=1)
        names = ['char_input'] + ['context%d_input' % (k + 1) for k in range(len(ctx))]
        self.model.inputs = [_Named(n + ':0') for n in names]
        self.status = 2

    @classmethod
    def from_weights(cls, wd, mapping_array, length, stateful=True, incremental=False, logger=None, device=None):
        """Ready-to-use Rater from in-memory weights (tests, benchmarks; no model file involved)."""
        r = cls(logger=logger, device=device)
        r.width = int(np.asarray(wd['char_embedding']).shape[1])
        r.depth = sum(1 for k in wd if k.startswith('lstm_'))
        r.length = length
        r.set_mapping(mapping_array)
        r.stateful, r.incremental = stateful, incremental
        r.configure()
        r.set_weights(wd)
        return r

    def save(self, filename):
        """Write weights + /config in the reference's file layout (reference :918-945)."""
        assert self.status > 1
        write_model_file(filename, self._host_weights, self.width, self.depth, self.length, self.stateful,
                         self.variable_length, self.mapping, self.voc_size, self.history)

    def reset_states(self):
        """Zero the implicit state of the stateful model (Keras ``model.reset_states()``, :475,:555)."""
        self._native.state_reset([0])

    def underspecify_contexts(self):
        """Zero (= unknown) for every context variable (reference :181-185)."""
        return [0] * len(self.context_shapes)

    def print_charset(self):
        for i, c in self.mapping[1].items():
            print('%d: "%s"' % (i, c))
            norm = unicodedata.normalize('NFC', c)
            if c != norm:
                self.logger.warning('mapped character "%s" (%d) should have been normalized to "%s", which is %s mapped',
                                    c, i, norm, 'also' if norm in self.mapping[0] else 'not')

    def print_history(self):
        for k, v in self.history.items():
            print("%s: %s" % (k, v))

    def train(self, *args, **kwargs):
        raise NotImplementedError('training is outside the B200 rating path; train with the reference and load the file')

    def remove_from_mapping(self, char=None, idx=None):
        """Prune one character from the vocabulary (reference :416-460, CLI ``prune-charset``): its row leaves the
        (tied) embedding, every higher index moves down by one, and the device model is rebuilt from the pruned
        weights.  Done as array surgery on the stored code-point array (index = vocabulary id) and the embedding
        matrix, not entry by entry.  Returns False (after logging) if there is nothing to remove."""
        assert self.status > 1
        assert self.voc_size > 0
        c_i, i_c = self.mapping
        if char:
            idx = c_i.get(char)
            if idx is None:
                self.logger.error('unmapped character "%s" cannot be removed', char)
                return False
        elif idx:
            char = i_c.get(idx)
            if char is None:
                self.logger.error('unmapped index "%d" cannot be removed', idx)
                return False
        else:
            return False
        embedding = np.asarray(self._host_weights['char_embedding'], dtype=np.float32)
        self.logger.warning('pruning character "%s" [%d] with norm %f', char, idx, np.linalg.norm(embedding[idx]))
        codepoints = np.zeros(embedding.shape[0], dtype=np.uint32)
        known = np.fromiter(i_c.keys(), dtype=np.int64, count=len(i_c))
        codepoints[known] = [ord(i_c[int(i)]) for i in known]
        wd = dict(self._host_weights)
        wd['char_embedding'] = np.delete(embedding, idx, axis=0)
        self.set_mapping(np.delete(codepoints, idx))        # rebuilds both dictionaries and voc_size
        self._lut = self._lut_src = None                    # the char -> index lookup table is stale now
        self.configure()
        self.set_weights(wd)
        return True

    # ------------------------------------------------------------------ vocabulary
    def _indices(self, text):
        """(int32 indices, bool mapped) of every character of `text` (one table lookup, no Python loop)."""
        if self._lut_src is not self.mapping[0] or self._lut is None:
            top = max([ord(c) for c in self.mapping[0]] + [0x7f]) + 1
            lut = np.zeros(top + 1, dtype=np.int32)
            known = np.zeros(top + 1, dtype=bool)
            for c, i in self.mapping[0].items():
                lut[ord(c)] = i
                known[ord(c)] = True
            self._lut, self._known, self._lut_src = lut, known, self.mapping[0]
        cps = np.frombuffer(text.encode('utf-32-le', 'surrogatepass'), dtype='<u4')
        idx = self._lut[np.minimum(cps, np.uint32(len(self._lut) - 1))]
        return idx, idx != 0                    # index 0 is never given to a mapped character

    def _context_row(self, context):
        if not context:
            context = self.underspecify_contexts()
        context = [int(c) for c in context]
        if len(context) != len(self.context_shapes):
            raise ValueError('model has %d context variables, got %d' % (len(self.context_shapes), len(context)))
        for c, (rows, _) in zip(context, self.context_shapes):
            if c < 0 or c >= rows:
                raise ValueError('context value %d outside embedding table of %d rows' % (c, rows))
        return context

    def _context_rows(self, contexts, n):
        """[n][number of context variables] int32; same checks as `_context_row`, done on the whole array."""
        if contexts is None:
            contexts = [None] * n
        if len(contexts) != n:
            raise ValueError('%d texts but %d contexts' % (n, len(contexts)))
        default = None
        try:
            rows = contexts
            if not all(contexts):                   # None or [] for a text: every variable unknown (:181-185)
                rows = []
                for c in contexts:
                    if not c:
                        if default is None:
                            default = [int(x) for x in self.underspecify_contexts()]
                        c = default
                    rows.append(c)
            arr = np.asarray(rows, dtype=np.int64)
            ok = arr.shape == (n, len(self.context_shapes))
            if ok and n:
                limit = np.array([r for r, _ in self.context_shapes], dtype=np.int64)
                ok = bool(((arr >= 0) & (arr < limit)).all())
        except (ValueError, TypeError):
            ok = False
        if not ok:                              # let the per-row check name the offending value
            arr = np.array([self._context_row(c) for c in contexts], dtype=np.int64).reshape(n, len(self.context_shapes))
        return arr.astype(np.int32)

    # ------------------------------------------------------------------ stateful rating
    def rate(self, text, context=None):
        """Probability of every character of `text` given its left context (reference :493-529).

        One kernel launch streams the whole string.  State is carried over from the previous call
        and NOT reset; like the reference's window generator (:1030-1057, :1096-1102) the last,
        partially filled window of `length` is executed including its zero padding (char 0,
        context 0), so that the carried state is the same as the reference's.
        """
        assert self.status > 1
        assert self.incremental is False
        if not self.stateful:
            # The reference's stateless mode scores every position from a zero state over a sliding window of at most
            # `length` characters (:1060-1094) -- a different quantity from the stateful recurrence, used only by models
            # trained stateless, and outside the B200 path: refuse rather than return other numbers.
            raise NotImplementedError('stateless (windowed) rating is not part of the B200 rating path; load the model stateful')
        context = self._context_row(context)
        text = unicodedata.normalize('NFC', text)
        size = len(text)
        if size <= 1:
            return [1.0] * size          # the reference fails inside Keras here
        steps = self.length if self.stateful else 1
        idx, known = self._indices(text)
        n_in = size - 1
        n_pad = -(-n_in // steps) * steps if self.stateful else n_in
        for t in np.flatnonzero(~known[:n_in]):
            self.logger.error('unmapped character "%s" at input position %d', text[t], int(t) % steps)
        for t in np.flatnonzero(~known[1:]):
            self.logger.error('unmapped character "%s" at output position %d', text[t + 1], int(t) % steps)
        chars = np.zeros(n_pad, dtype=np.int32)
        chars[:n_in] = idx[:n_in]
        targets = np.zeros(n_pad, dtype=np.int32)
        targets[:n_in] = idx[1:]
        logp = np.empty(n_pad, dtype=np.float32)
        if not self.stateful:
            self._native.state_reset([0])
        self._native.rate([0], [0, n_pad], chars, targets, [n_in], [context], logp)
        return [1.0] + list(np.exp(logp[:n_in]))

    def rate2(self, text, context=None):
        """Rate a string from a reset state, no padding (reference :531-576).

        Returns ([(char, prob), ...], perplexity).  Same launch as `rate`; the reference needed
        one device call per character here.
        """
        assert self.status > 1
        assert self.incremental is False
        assert self.stateful
        context = self._context_row(context)
        text = unicodedata.normalize('NFC', text)
        size = len(text)
        if size == 0:
            return [], 1.0
        idx, known = self._indices(text)
        for t in np.flatnonzero(~known):
            self.logger.error('unmapped character "%s" at input position %d', text[t], int(t))
        self._native.state_reset([0])
        probs = np.ones(size, dtype=np.float32)
        if size > 1:
            chars = np.ascontiguousarray(idx[:-1], dtype=np.int32)
            targets = np.ascontiguousarray(idx[1:], dtype=np.int32)
            logp = np.empty(size - 1, dtype=np.float32)
            self._native.rate([0], [0, size - 1], chars, targets, None, [context], logp)
            probs[1:] = np.exp(logp)
        entropy = 0.0
        for p in probs[1:]:
            entropy -= math.log(max(float(p), 1e-99), 2)
        result = [(c, 1.0 if i == 0 else float(p)) for i, (c, p) in enumerate(zip(text, probs))]
        return result, pow(2.0, entropy / size)

    def rate_many(self, texts, contexts=None, return_logp=False):
        """Rate independent texts concurrently, each from a zero state (no carry, no padded tail).

        This is the throughput form of `rate`: what a caller gets from a fresh/reset ``Rater`` per
        document, page or line, with all streams advancing together in one launch per
        batch.  Returns one float32 array of probabilities per text (first entry 1.0).
        """
        assert self.status > 1
        assert self.incremental is False
        n = len(texts)
        ctx_rows = self._context_rows(contexts, n)
        norm = [t if t.isascii() else unicodedata.normalize('NFC', t) for t in texts]
        nchar = np.fromiter(map(len, norm), dtype=np.int64, count=n)
        report = self.logger.isEnabledFor(logging.ERROR)    # unmapped characters are reported like the reference does (:1116)
        lens = np.maximum(nchar - 1, 0)
        order = np.argsort(-lens, kind='stable')            # longest first: batches of similar length
        total = int(lens.sum())
        # Many short texts (lines, words): ONE table lookup for the characters of all texts, and the segments are
        # gathered / scattered with index arithmetic -- no per-text Python work.  Few long texts (documents): every
        # segment maps its slices itself, so that this work overlaps the device (a lookup over all characters up
        # front would not).
        gather = total < n * self.GATHER_MAX_MEAN_LENGTH
        cstart = np.zeros(n + 1, dtype=np.int64)
        np.cumsum(nchar, out=cstart[1:])
        if gather:
            # text i occupies [cstart[i], cstart[i+1]) of idx_all / out_all
            idx_all, known_all = self._indices(''.join(norm))
            if report and not known_all.all():
                bad = np.flatnonzero(~known_all)
                own = np.searchsorted(cstart, bad, side='right') - 1
                for i, pos in zip(own.tolist(), (bad - cstart[own]).tolist()):
                    self.logger.error('unmapped character "%s" at input position %d', norm[i][pos], pos)
            out_all = np.empty(int(cstart[-1]), dtype=np.float32)
            out_all[cstart[:-1][nchar > 0]] = 0.0 if return_logp else 1.0
            edges = cstart.tolist()
            out = _Scores(out_all[a:b] for a, b in zip(edges, edges[1:]))
            out.flat = out_all                      # the same scores as ONE array in input order (rate_sharded sends it as is)
        else:
            out = [np.empty(len(t), dtype=np.float32) for t in norm]
            for o in out:
                o[:1] = 0.0 if return_logp else 1.0
        if total == 0:
            return out
        # The work is cut into segments and a worker thread drives the device: while segment k runs, the host maps
        # the characters of segment k+1 and converts the scores of segment k-1.  Long texts are cut along time (every
        # stream contributes its share to each segment, the state is carried in the pool slots); many short texts
        # are cut into groups of whole streams (at multiples of the launch width, so that launches stay full) -- cut
        # along time they would pay the pipeline fill of a launch once per segment.
        nseg = int(min(self.MAX_SEGMENTS, max(1, total // self.MIN_SEGMENT_POSITIONS)))
        lens_o, cstart_o = lens[order], cstart[:-1][order]
        if gather:
            cum = np.cumsum(lens_o)
            cuts = [0]
            for k in range(1, nseg):
                c = int(np.searchsorted(cum, total * k // nseg, side='left')) + 1
                c = min(n, -(-c // self.LAUNCH_STREAMS) * self.LAUNCH_STREAMS)
                if c > cuts[-1]:
                    cuts.append(c)
            if cuts[-1] < n:
                cuts.append(n)
            nseg = len(cuts) - 1
            seg_tot = [int(lens_o[cuts[k]:cuts[k + 1]].sum()) for k in range(nseg)]
        else:
            bounds = [(lens_o * k) // nseg for k in range(nseg + 1)]           # [nseg+1][n] position boundaries
            seg_tot = [int((bounds[k + 1] - bounds[k]).sum()) for k in range(nseg)]
        bufs = self._pinned(max(seg_tot), 2 if nseg > 1 else 1)
        ctx_rows = ctx_rows[order]
        slots = self._pool.acquire(n)

        def prepare(k):
            chars, targets, logp = (a[:seg_tot[k]] for a in bufs[k % len(bufs)])
            if gather:
                a, b = cuts[k], cuts[k + 1]
                counts = lens_o[a:b]
                offsets = np.zeros(b - a + 1, dtype=np.int64)
                np.cumsum(counts, out=offsets[1:])
                # position p of stream j is character cstart_o[j] + (p - offsets[j]) of idx_all
                src = np.repeat(cstart_o[a:b] - offsets[:-1], counts)
                src += np.arange(seg_tot[k], dtype=np.int64)
                np.take(idx_all, src, out=chars, mode='clip')
                src += 1
                np.take(idx_all, src, out=targets, mode='clip')
                return slots[a:b], offsets, ctx_rows[a:b], logp, src            # src now addresses the scored characters
            counts = bounds[k + 1] - bounds[k]
            offsets = np.zeros(n + 1, dtype=np.int64)
            np.cumsum(counts, out=offsets[1:])
            for j, i in enumerate(order):
                lo, hi = int(bounds[k][j]), int(bounds[k + 1][j])
                if hi == lo:
                    if report and k == 0 and len(norm[i]) == 1 and not self._indices(norm[i])[1][0]:
                        self.logger.error('unmapped character "%s" at input position %d', norm[i], 0)
                    continue
                idx, known = self._indices(norm[i][lo:hi + 1])
                if k > 0:
                    known = known[1:]               # the segment's first character was reported with the previous one
                    base = lo + 1
                else:
                    base = lo
                if report:
                    for t in np.flatnonzero(~known).tolist():
                        self.logger.error('unmapped character "%s" at input position %d', norm[i][base + t], base + t)
                chars[offsets[j]:offsets[j + 1]] = idx[:-1]
                targets[offsets[j]:offsets[j + 1]] = idx[1:]
            return slots, offsets, ctx_rows, logp, None

        def finish(k, offsets, logp, dst):
            if gather:
                out_all[dst] = logp if return_logp else np.exp(logp)
                return
            for j, i in enumerate(order):
                lo, hi = int(bounds[k][j]), int(bounds[k + 1][j])
                if hi > lo:
                    seg = logp[offsets[j]:offsets[j + 1]]
                    if return_logp:
                        out[i][1 + lo:1 + hi] = seg
                    else:
                        np.exp(seg, out=out[i][1 + lo:1 + hi])

        try:
            self._native.state_reset(slots)
            if nseg == 1:
                sl, offsets, cx, logp, dst = prepare(0)
                chars, targets = (a[:seg_tot[0]] for a in bufs[0][:2])
                self._native.rate(sl, offsets, chars, targets, None, cx, logp)
                finish(0, offsets, logp, dst)
            else:
                if self._worker is None:
                    from concurrent.futures import ThreadPoolExecutor
                    self._worker = ThreadPoolExecutor(max_workers=1, thread_name_prefix='klm-rate')
                pending = None                      # (future, k, offsets, logp) of the segment on the device
                try:
                    for k in range(nseg):
                        sl, offsets, cx, logp, dst = prepare(k)
                        chars, targets = (a[:seg_tot[k]] for a in bufs[k % len(bufs)][:2])
                        if pending is not None:
                            pending[0].result()
                        fut = self._worker.submit(self._native.rate, sl, offsets, chars, targets, None, cx, logp)
                        if pending is not None:
                            finish(*pending[1:])
                        pending = (fut, k, offsets, logp, dst)
                    pending[0].result()
                    finish(*pending[1:])
                    pending = None
                finally:
                    if pending is not None:         # an exception above: never leave the device call running
                        try:
                            pending[0].result()
                        except Exception:           # pylint: disable=broad-except
                            pass
        finally:
            self._pool.release_many(slots)
        return out

    GATHER_MAX_MEAN_LENGTH = 4096         # rate_many: one lookup + index arithmetic below this mean text length
    LAUNCH_STREAMS = 64                   # rate_many: streams per launch of the sequence kernels
    MAX_SEGMENTS = 16                     # rate_many: segments for host/device overlap ...
    MIN_SEGMENT_POSITIONS = 250000        # ... each at least this many positions (a launch costs ~50 us)

    def _pinned(self, total, sets=1):
        """`sets` groups of page-locked (chars, targets, scores) buffers for `total` positions each
        (grow-only, reused)."""
        if self._pin is None or len(self._pin) < sets or len(self._pin[0][0]) < total:
            cap = total + total // 8 + 1024
            self._pin = [(_native.pinned_empty(cap, np.int32), _native.pinned_empty(cap, np.int32),
                          _native.pinned_empty(cap, np.float32)) for _ in range(sets)]
        return self._pin[:sets]

    def test(self, test_data):
        """Perplexity over text files (reference :462-491, :977-1002).

        As in the reference: state reset once (not per file), context = ceil(year/10) from file
        names `author_title_year`, loss = mean over windows of the mean over all `length`
        positions of the window (padding positions count with zero loss), result exp(loss).
        """
        assert self.status > 1
        assert self.incremental is False
        if self.stateful:
            self.reset_states()
        steps = self.length if self.stateful else 1
        losses = []
        for file in test_data:
            file.seek(0)
            parts = os.path.basename(file.name).split('.')[0].split('_')
            year = math.ceil(int(parts[2]) / 10) if len(parts) == 3 else 0
            text = file.read()
            if isinstance(text, bytes):
                text = text.decode('utf-8')
            text = unicodedata.normalize('NFC', text)
            if len(text) <= 1:
                continue
            probs = np.asarray(self.rate(text, [year])[1:], dtype=np.float64)
            ce = -np.log(np.clip(probs, 1e-7, 1 - 1e-7))
            for k in range(0, len(ce), steps):
                losses.append(ce[k:k + steps].sum() / steps)
        return math.exp(float(np.mean(losses)))

    # ------------------------------------------------------------------ incremental step
    def _slots_for(self, initial_states):
        """Pool slots of the given states (-1 = zero state); uploads caller-made array lists."""
        slots = np.empty(len(initial_states), dtype=np.int32)
        temps = []
        for i, st in enumerate(initial_states):
            if isinstance(st, DeviceState):
                slots[i] = st.slot
            elif st is None or len(st) == 0:
                slots[i] = -1
            else:                                   # list of 2*depth arrays made by the caller
                arr = np.stack([np.asarray(a, dtype=np.float32).reshape(-1) for a in st])
                if not arr.any():
                    slots[i] = -1
                    continue
                tmp = DeviceState(self._pool, self._pool.acquire(1)[0])
                self._native.state_from_host([tmp.slot], arr[None])
                temps.append(tmp)
                slots[i] = tmp.slot
        return slots, temps

    def _advance(self, indices, initial_states, context, targets=None):
        """One timestep for n hypotheses on the device.

        indices: input character index per hypothesis; targets: None -> full softmax rows (n, V),
        else the natural-log probability of targets[i] only (what beam search needs).
        Returns (scores, [DeviceState]*n).
        """
        n = len(indices)
        slots_in, temps = self._slots_for(initial_states)
        out_slots = self._pool.acquire(n)
        new_states = [DeviceState(self._pool, s) for s in out_slots]
        ctx = np.tile(np.asarray(context, dtype=np.int32), (n, 1)) if len(context) else None
        if targets is None:
            out = np.empty((n, self._native.voc_size), dtype=np.float32)
        else:
            out = np.empty(n, dtype=np.float32)
        self._native.step(indices, ctx, slots_in, out_slots, targets, out)
        self.device_steps += 1
        self.hypothesis_steps += n
        del temps
        return out, new_states

    def _advance_topk(self, indices, initial_states, context, k):
        """One timestep for n hypotheses; only the k most probable next characters of each come back
        (indices [n, k], probabilities [n, k], ascending by probability) + the new states."""
        n = len(indices)
        slots_in, temps = self._slots_for(initial_states)
        out_slots = self._pool.acquire(n)
        new_states = [DeviceState(self._pool, s) for s in out_slots]
        ctx = np.tile(np.asarray(context, dtype=np.int32), (n, 1)) if len(context) else None
        tops, probs = self._native.step_topk(indices, ctx, slots_in, out_slots, k)
        self.device_steps += 1
        self.hypothesis_steps += n
        del temps
        return tops, probs, new_states

    def predict(self, candidates, initial_states, context=None):
        """Next-character distributions and new states for n hypotheses (reference :578-639).

        `initial_states[i]` may be None (zero state), a state returned by an earlier call, or a
        list of 2*depth arrays.  Returns (list of n arrays (V,), list of n states).  As in the
        reference, None entries of `initial_states` are replaced in place by zero states (:625-627).
        """
        assert self.status > 1
        assert self.stateful is False
        assert self.incremental is True
        assert len(candidates) == len(initial_states), \
            "number of inputs (%d) and number of states (%d) inconsistent" % (len(candidates), len(initial_states))
        context = self._context_row(context)
        n = len(candidates)
        if n == 0:
            return [], []
        c_i = self.mapping[0]
        indices = np.fromiter((c_i.get(c, 0) for c in candidates), dtype=np.int32, count=n)
        for i, st in enumerate(initial_states):
            if not st:
                initial_states[i] = [np.zeros((self.width), dtype=float) for _ in range(self.depth * 2)]
        probs, states = self._advance(indices, initial_states, context)
        return [probs[i, :] for i in range(n)], states

    def generate(self, prefix, length, context=None, variants=1):
        """Continue `prefix` by `length` characters with a beam search (reference :642-709)."""
        assert self.status > 1
        assert self.stateful is False
        assert self.incremental is True
        context = self._context_row(context)
        states = [None]
        for char in prefix[:-1]:
            _, states = self.predict([char], states, context=context)
        fringe = [Node(state=states[0], value=prefix[-1], cost=0.0)]
        c_i, i_c = self.mapping
        for _ in range(length):
            current = fringe
            # the ten best continuations of every hypothesis are picked on the device (klm_step_topk): 10 (index,
            # probability) pairs per hypothesis come back instead of a softmax row
            indices = np.fromiter((c_i.get(n.value, 0) for n in current), dtype=np.int32, count=len(current))
            tops, probs, states = self._advance_topk(indices, [n.state for n in current], context, min(10, self.voc_size))
            fringe = []
            for j, node in enumerate(current):
                first = np.searchsorted(probs[j], 0.004)           # ascending: keep what reaches the absolute threshold
                for v, cost in zip(tops[j, first:], -np.log(probs[j, first:])):
                    if v not in i_c:
                        continue
                    bisect.insort_left(fringe, Node(parent=node, state=states[j], value=i_c[v], cost=cost))
            fringe = fringe[:256]
        return [''.join(n.value for n in best.to_sequence()) for best in fringe[0:variants]]

    # ------------------------------------------------------------------ lattice beam search
    def rate_best(self, graph, start_node, end_node, start_traceback=None, context=None, lm_weight=0.5,
                  beam_width=10, beam_clustering_dist=0):
        """Decode the best path through a lattice of alternatives (reference :712-859).

        `graph`: networkx.DiGraph whose edges carry `element` and `alternatives` (objects with
        `.Unicode`, `.conf`, `.index`).  Control flow, pruning constants and orderings are the
        reference's; per character step the device is asked only for the
        log-probability of the one character each candidate wants next (fused projection +
        log-softmax + pick), and states never leave the device.
        Returns (path of the PREVIOUS graph [(element, alternative, score)], entropy,
        (traceback beam, start node)).
        """
        import networkx as nx
        assert self.status > 1
        assert self.stateful is False
        assert self.incremental is True
        context = self._context_row(context)
        if not start_traceback:
            root = Node(state=None, value='\n', cost=0.0)
            start_traceback = ([root], root)
        graph.nodes[start_node]['traceback'] = start_traceback[0]
        c_i = self.mapping[0]
        # edges in the order the reference visits them (:763-770): topological order of the heads,
        # every in-edge whose tail has been reached (the reached-list may hold duplicates)
        reached = [start_node]
        order = []
        for head in nx.topological_sort(graph):
            for tail, _ in graph.in_edges([head]):
                if tail in reached:
                    order.append((tail, head))
                    reached.append(head)
        out_node = None
        out_attrs = None
        if self.native_beam and self.native_lattice and not beam_clustering_dist and self._is_chain(graph, order, start_node):
            # a linear lattice (what a PAGE gives): all edges in ONE native call, tree nodes only for what is looked at later
            out_node = order[-1][1]
            out_attrs = graph.nodes[out_node]
            out_attrs['traceback'] = self._beam_lattice_native(graph, order, start_traceback[0], context, lm_weight, beam_width)
            order = []
        for in_node, out_node in order:
            edge = graph.edges[in_node, out_node]
            element, alternatives = edge['element'], edge['alternatives']
            in_attrs, out_attrs = graph.nodes[in_node], graph.nodes[out_node]
            assert 'traceback' in in_attrs, "graph not topologically connected at node %s" % str(in_node)
            if self.native_beam and not beam_clustering_dist and 'traceback' not in out_attrs:
                out_attrs['traceback'] = self._beam_edge_native(in_attrs['traceback'], element, alternatives, context,
                                                                lm_weight, beam_width)
                continue
            done = out_attrs['traceback'] if 'traceback' in out_attrs else []
            waiting = [Node(parent=hyp, state=hyp.state, value="", cost=0.0, extras=(element, alt))
                       for hyp in in_attrs['traceback'] for alt in alternatives]
            reported = {}
            rounds = 3 * max(len(alt.Unicode) for alt in alternatives)
            for _ in range(rounds):
                batch = []
                while waiting:                               # most expensive first
                    cand = waiting.pop()
                    if cand.value == cand.extras[1].Unicode:
                        if beam_clustering_dist and self._history_clustering(cand, done, beam_clustering_dist):
                            continue
                        bisect.insort_left(done, cand)
                    else:
                        bisect.insort_left(batch, cand)
                    if len(batch) >= self.batch_size:
                        break
                if not batch:
                    break
                if done and batch[0].cum_cost >= done[0].cum_cost + 15:
                    break
                n = len(batch)
                feed = np.empty(n, dtype=np.int32)
                want = np.empty(n, dtype=np.int32)
                nxt = []
                for i, cand in enumerate(batch):
                    prev = cand.value[-1] if cand.value else cand.parent.value[-1]
                    feed[i] = c_i.get(prev, 0)
                    alt = cand.extras[1]
                    ch = alt.Unicode[len(cand.value)]
                    if ch not in c_i:
                        seen = reported.setdefault(alt.index, [])
                        if ch not in seen:
                            self.logger.error('unmapped character "%s" at input alternative %d of element %s',
                                              ch, alt.index or i, element.id if element else "space")
                            seen.append(ch)
                    want[i] = c_i.get(ch, 0)
                    nxt.append(ch)
                logp, states = self._advance(feed, [cand.state for cand in batch], context, targets=want)
                probs = np.exp(logp)                          # float32, like the softmax output it replaces
                for i, cand in enumerate(batch):
                    conf = cand.extras[1].conf
                    cand.cum_cost += (-math.log(max(float(probs[i]), 1e-99), 2) * lm_weight +
                                      -math.log(max(conf, 1e-99), 2) * (1. - lm_weight))
                    cand.value += nxt[i]
                    cand.state = states[i]
                    if waiting and cand.cum_cost >= waiting[0].cum_cost + 2.5:
                        continue
                    bisect.insort_left(waiting, cand)
                waiting = waiting[:rounds * self.batch_size]
            out_attrs['traceback'] = done[:beam_width]
        assert out_node == end_node, 'graph must have a path from start_node to end_node'
        assert out_attrs is not None and 'traceback' in out_attrs, 'graph has no traceback at end_node'
        return self.next_path(out_attrs['traceback'], start_traceback)

    def _beam_edge_native(self, beam, element, alternatives, context, lm_weight, beam_width):
        """One lattice edge through ``klm_beam_edge``: the reference's per-edge loop (:789-852) run natively.

        Only the finished candidates that make it into the beam become ``Node`` objects; their states stay in
        the device pool.  Unmapped characters are reported once per alternative up front.
        """
        c_i = self.mapping[0]
        n_hyp, n_alt = len(beam), len(alternatives)
        offsets = np.zeros(n_alt + 1, dtype=np.int32)
        chars = []
        for a, alt in enumerate(alternatives):
            seen = []
            for ch in alt.Unicode:
                if ch not in c_i and ch not in seen:
                    self.logger.error('unmapped character "%s" at input alternative %d of element %s',
                                      ch, alt.index or 0, element.id if element else "space")
                    seen.append(ch)
                chars.append(c_i.get(ch, 0))
            offsets[a + 1] = len(chars)
        if n_hyp == 0:
            return []
        slots, temps = self._slots_for([hyp.state for hyp in beam])
        last = np.fromiter((c_i.get(hyp.value[-1], 0) for hyp in beam), dtype=np.int32, count=n_hyp)
        cum = np.fromiter((hyp.cum_cost for hyp in beam), dtype=np.float64, count=n_hyp)
        conf = np.fromiter((alt.conf for alt in alternatives), dtype=np.float64, count=n_alt)
        scratch = self._pool.acquire(n_hyp * n_alt + self.batch_size)
        try:
            oh, oa, oc, osl, steps, hsteps = self._native.beam_edge(
                slots, last, cum, offsets, np.asarray(chars, dtype=np.int32), conf, context, lm_weight, self.batch_size,
                scratch, beam_width)
        except Exception:
            self._pool.release_many(scratch)
            raise
        self.device_steps += steps
        self.hypothesis_steps += hsteps
        kept = set(osl[osl >= 0].tolist())
        self._pool.release_many([sl for sl in scratch if sl not in kept])
        del temps
        out = []
        for h, a, cost, sl in zip(oh, oa, oc, osl):
            hyp, alt = beam[int(h)], alternatives[int(a)]
            state = DeviceState(self._pool, int(sl)) if sl >= 0 else hyp.state
            node = Node(parent=hyp, state=state, value=alt.Unicode, cost=0.0, extras=(element, alt))
            node.cum_cost = float(cost)
            out.append(node)
        return out

    @staticmethod
    def _is_chain(graph, order, start_node):
        """True if the edges in search order form one path from `start_node` through nodes that carry no beam yet."""
        if not order or order[0][0] != start_node:
            return False
        for k, (tail, head) in enumerate(order):
            if k and tail != order[k - 1][1]:
                return False
            if 'traceback' in graph.nodes[head] or not graph.edges[tail, head]['alternatives']:
                return False
        return len(set(head for _, head in order)) == len(order)

    def _beam_lattice_native(self, graph, order, beam, context, lm_weight, beam_width):
        """All edges of a linear lattice through ``klm_beam_lattice`` (the per-edge loop of reference :789-852, edge after
        edge, natively).  Returns the beam at the last node; ``Node`` objects are made only for the final beam and its
        ancestors (everything `next_path` can reach), the rest of the back-pointer table is dropped."""
        c_i = self.mapping[0]
        edges = [graph.edges[tail, head] for tail, head in order]
        edge_first = np.zeros(len(edges) + 1, dtype=np.int32)
        offsets, chars, conf = [0], [], []
        max_alts = 1
        for e, edge in enumerate(edges):
            element, alternatives = edge['element'], edge['alternatives']
            max_alts = max(max_alts, len(alternatives))
            for alt in alternatives:
                seen = []
                for ch in alt.Unicode:
                    if ch not in c_i and ch not in seen:
                        self.logger.error('unmapped character "%s" at input alternative %d of element %s',
                                          ch, alt.index or 0, element.id if element else "space")
                        seen.append(ch)
                    chars.append(c_i.get(ch, 0))
                offsets.append(len(chars))
                conf.append(alt.conf)
            edge_first[e + 1] = len(conf)
        n_hyp = len(beam)
        if n_hyp == 0:
            return []
        slots, temps = self._slots_for([hyp.state for hyp in beam])
        last = np.fromiter((c_i.get(hyp.value[-1], 0) for hyp in beam), dtype=np.int32, count=n_hyp)
        cum = np.fromiter((hyp.cum_cost for hyp in beam), dtype=np.float64, count=n_hyp)
        width = max(beam_width, n_hyp)
        scratch = self._pool.acquire(width * (max_alts + 1) + self.batch_size + width)
        try:
            first, oh, oa, oc, fslot, steps, hsteps = self._native.beam_lattice(
                slots, last, cum, edge_first, np.asarray(offsets, dtype=np.int32), np.asarray(chars, dtype=np.int32),
                np.asarray(conf, dtype=np.float64), context, lm_weight, self.batch_size, beam_width, scratch)
        except Exception:
            self._pool.release_many(scratch)
            raise
        self.device_steps += steps
        self.hypothesis_steps += hsteps
        n_final = int(first[-1] - first[-2])
        kept = set(int(s) for s in fslot[:n_final] if s >= 0)
        self._pool.release_many([sl for sl in scratch if sl not in kept])
        del temps
        made = {}

        def node_at(e, row):
            """Tree node of kept candidate `row` of edge `e` (memoised; ancestors first, iteratively)."""
            todo = []
            while e >= 0 and (e, row) not in made:
                todo.append((e, row))
                row = int(oh[first[e] + row])
                e -= 1
            parent = beam[row] if e < 0 else made[(e, row)]
            for e2, row2 in reversed(todo):
                k = int(first[e2]) + row2
                edge = edges[e2]
                alt = edge['alternatives'][int(oa[k])]
                node = Node(parent=parent, state=parent.state if not alt.Unicode else None, value=alt.Unicode, cost=0.0,
                            extras=(edge['element'], alt))
                node.cum_cost = float(oc[k])
                made[(e2, row2)] = node
                parent = node
            return parent

        out = []
        last_edge = len(edges) - 1
        for row in range(n_final):
            node = node_at(last_edge, row)
            sl = int(fslot[row])
            if sl >= 0:
                node.state = DeviceState(self._pool, sl)
            elif sl <= -2:
                node.state = beam[-2 - sl].state
            out.append(node)
        return out

    def next_path(self, beam, traceback):
        """Decide the previous graph's path from the current beam and cut histories there (reference :862-885)."""
        prev_beam, prev_start = traceback
        best_path = beam[0].to_sequence(stop_at=prev_beam)
        start_node = best_path[-1]
        result = []
        for node in best_path:
            if not node.extras:
                continue
            element, alt = node.extras
            before = node.parent.cum_cost if node.parent else prev_start.cum_cost
            result.append((element, alt, pow(2.0, -(node.cum_cost - before) / len(alt.Unicode))))
        next_beam = []
        for hyp in beam:
            if hyp.to_sequence(stop_at=[start_node]):
                hyp.cut_at(start_node)
                bisect.insort_left(next_beam, hyp)
        return result, start_node.cum_cost - prev_start.cum_cost, (next_beam, start_node)

    def _history_clustering(self, cand, done, distance=5):
        """True if `cand` duplicates a cheaper finished hypothesis with (nearly) the same history (reference :887-916).

        Same value on this edge and |state difference| < distance for the first `depth` arrays of
        the h1,c1,h2,.. list; a more expensive duplicate in `done` is evicted instead.
        """
        for other in done:
            if cand.value != other.value:
                continue
            a, b = cand.state, other.state
            if isinstance(a, DeviceState) and isinstance(b, DeviceState):
                near = bool((self._native.state_l2_distance([a.slot], [b.slot], self.depth)[0] < distance).all())
            else:
                near = all(np.linalg.norm(np.asarray(a[k]) - np.asarray(b[k])) < distance for k in range(self.depth))
            if near:
                if other.cum_cost < cand.cum_cost:
                    return True
                done.remove(other)
                break
        return False


# ---------------------------------------------------------------------- helpers
class _Named(object):
    def __init__(self, name):
        self.name = name


def _as_str(v):
    if isinstance(v, bytes):
        return v.decode('utf-8')
    if isinstance(v, np.ndarray):
        return _as_str(v.item() if v.shape == () else v.tobytes())
    return str(v)


def _layer_weights(f, layer):
    """Weight arrays of one layer group of a Keras weights file, in `weight_names` order."""
    g = f[layer]
    names = [(_as_str(n)) for n in g.attrs.get('weight_names', [])]
    return [g[n][()] for n in names]


def read_keras_weights(f, depth):
    """{'char_embedding': E, 'context%d_embedding': C, 'lstm_%d': (K, R, b)} from an open model file.

    Keras' own loader is topological (k-th weighted layer of the file feeds the k-th of the
    model); with the reference's fixed layer names, loading by name is equivalent.
    """
    wd = {'char_embedding': _layer_weights(f, 'char_embedding')[0]}
    k = 1
    while 'context%d_embedding' % k in f:
        wd['context%d_embedding' % k] = _layer_weights(f, 'context%d_embedding' % k)[0]
        k += 1
    for l in range(depth):
        name = 'lstm_%d' % (l + 1)
        if name not in f:
            raise ValueError('model file has no layer %s' % name)
        ws = _layer_weights(f, name)
        if len(ws) != 3:
            raise ValueError('layer %s has %d weight arrays, expected 3' % (name, len(ws)))
        wd[name] = tuple(ws)
    return wd


def convert_cudnn_weights(K, R, b):
    """CuDNNLSTM-saved weights -> LSTM layout (Keras `_convert_rnn_weights`).

    Detected by a bias of 8*width (input bias | recurrent bias): per gate block the kernels
    are stored transposed; the two biases add up.
    """
    W = R.shape[0]
    if b.shape[0] != 8 * W:
        return K, R, b
    Kc = np.concatenate([K[:, g * W:(g + 1) * W].T.reshape(K[:, g * W:(g + 1) * W].shape, order='F')
                         for g in range(4)], axis=1)
    Rc = np.concatenate([R[:, g * W:(g + 1) * W].T for g in range(4)], axis=1)
    return np.ascontiguousarray(Kc), np.ascontiguousarray(Rc), b.reshape(2, -1).sum(axis=0)


def write_model_file(filename, wd, width, depth, length, stateful, variable_length, mapping, voc_size, history=None):
    """Model file in the reference's layout: Keras save_weights groups + /config (reference :918-945)."""
    nctx = sum(1 for k in wd if k.startswith('context'))
    layers = ['char_input'] + ['context%d_input' % (k + 1) for k in range(nctx)] + ['char_embedding'] + \
             ['context%d_embedding' % (k + 1) for k in range(nctx)] + ['concat_hidden_input']
    for l in range(depth):
        if l > 0:
            layers.append('dropout_%d' % l)
        layers.append('lstm_%d' % (l + 1))
    layers.append('char_output')
    with h5lite.File(filename, 'w') as f:
        f.attrs['layer_names'] = np.array([n.encode('utf-8') for n in layers])
        f.attrs['backend'] = np.bytes_(b'tensorflow')
        f.attrs['keras_version'] = np.bytes_(b'2.3.1')
        for name in layers:
            g = f.create_group(name)
            if name in wd:
                arrays = wd[name]
                if not isinstance(arrays, tuple):
                    arrays, wnames = (arrays,), ['embeddings:0']
                else:
                    wnames = ['kernel:0', 'recurrent_kernel:0', 'bias:0']
                full = ['%s/%s' % (name, w) for w in wnames]
                g.attrs['weight_names'] = np.array([n.encode('utf-8') for n in full])
                sub = g.create_group(name)
                for w, a in zip(wnames, arrays):
                    sub.create_dataset(w, data=np.asarray(a, dtype=np.float32))
            else:
                g.attrs['weight_names'] = np.zeros((0,), dtype='S1')
        cfg = f.create_group('config')
        cfg.create_dataset('history', data=json.dumps(history or {}))
        cfg.create_dataset('width', data=np.array(width))
        cfg.create_dataset('depth', data=np.array(depth))
        cfg.create_dataset('length', data=np.array(length))
        cfg.create_dataset('stateful', data=np.array(bool(stateful)))
        cfg.create_dataset('variable_length', data=np.array(bool(variable_length)))
        i_c = mapping[1]
        cfg.create_dataset('mapping', data=np.fromiter((ord(i_c[i]) if i in i_c else 0 for i in range(voc_size)),
                                                       dtype='uint32'))
