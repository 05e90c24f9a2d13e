"""CPU restatement of the reference ``Rater`` rating path on top of ``oracle.lstm_numpy``.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``): the checker for the CUDA
path and the timed CPU arm of ``bench.py``; never imported by the product.

Each method cites the part of /root/reference/ocrd_keraslm/lib/rating.py it
follows.  The restatement is pinned by ``tests/golden/*.json|npz``, which were
produced by running the reference's own rating.py (see
``tests/golden/make_golden.py``) on the same seeded models and inputs.
"""
import bisect
import logging
import math
import unicodedata

import numpy as np

from . import lstm_numpy as ln


class OracleNode(object):
    """Back-pointer tree node, ordered by projected cost (rating.py:1240-1318)."""

    def __init__(self, state, value, cost, parent=None, extras=None):
        self.value = value
        self.parent = parent
        self.state = state
        self.cum_cost = cost if parent is None else parent.cum_cost + cost      # :1262
        self.length = 1 if parent is None else parent.length + 1               # :1263
        self.extras = extras
        self._sequence = None

    # -- ordering and equality are all by projected cost (:1300-1318) --
    def pro_cost(self):
        remaining = len(self.extras[1].Unicode) - len(self.value) if self.extras else 0
        return self.cum_cost + 0.5 * remaining

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

    __hash__ = None

    def to_sequence(self, stop_at=None):
        """Root..self, or (first hit in stop_at)..self; cached on first call (:1269-1285)."""
        if not self._sequence:
            chain = []
            collecting = not stop_at
            node = self
            while node:       # NB truthiness of a Node is default (True)
                if stop_at and node in stop_at:     # membership by pro_cost equality (:1280)
                    collecting = True
                if collecting:
                    chain.append(node)
                node = node.parent
            chain.reverse()
            self._sequence = chain
        return self._sequence

    def cut_at(self, node):
        """Detach the ancestor whose parent *is* node; clears only own cache (:1287-1298)."""
        cur = self
        while cur:
            if cur.parent is node:
                cur.parent = None
                self._sequence = None
                return
            cur = cur.parent


class OracleRater(object):
    def __init__(self, weights, mapping, length, batch_size=128, logger=None):
        self.w = weights
        self.width, self.depth = weights.width, weights.depth
        self.voc_size = weights.voc_size
        self.mapping = mapping
        self.length = length
        self.batch_size = batch_size            # rating.py:49 (incremental mode keeps 128)
        self.logger = logger or logging.getLogger(__name__)
        self.states = ln.zero_states(weights, 1)
        self.device_calls = 0                    # number of model invocations (Session.run equivalents)

    # ------------------------------------------------------------------ helpers
    def reset_states(self):
        self.states = ln.zero_states(self.w, 1)

    def underspecify_contexts(self):              # rating.py:181-185
        return [0] * self.w.ncontexts

    def _index(self, char):
        return self.mapping[0].get(char, 0)

    # ------------------------------------------------------------------ rate
    def rate(self, text, context=None):
        """rating.py:493-529 + windowing of _gen_data :1030-1057,:1096-1102 and _vectorize :1123-1143.

        The network consumes text[0:T-1] in ceil((T-1)/L) windows of L; the last one is
        right-padded with (char 0, context 0) and the padded steps DO advance the carried
        state.  No reset between calls.
        """
        if not context:
            context = self.underspecify_contexts()
        text = unicodedata.normalize('NFC', text)
        T, L = len(text), self.length
        if T <= 1:
            return [1.0] * T                      # reference is undefined here
        n_windows = -(-(T - 1) // L)
        x = np.zeros(n_windows * L, dtype=np.int64)
        zs = [np.zeros(n_windows * L, dtype=np.int64) for _ in context]
        for t in range(T - 1):
            ch = text[t]
            if ch not in self.mapping[0]:
                self.logger.error('unmapped character "%s" at input position %d', ch, t % L)
            x[t] = self._index(ch)
            for z, c in zip(zs, context):
                z[t] = c
        targets = np.zeros(n_windows * L, dtype=np.int64)
        for t in range(1, T):
            if text[t] not in self.mapping[0]:                 # _vectorize also builds y (:1149-1151)
                self.logger.error('unmapped character "%s" at output position %d', text[t], (t - 1) % L)
            targets[t - 1] = self._index(text[t])
        out = np.empty(n_windows * L, dtype=self.w.E.dtype)
        for k in range(n_windows):                # one Session.run per window in the reference
            sl = slice(k * L, (k + 1) * L)
            o, self.states = ln.run_sequence(self.w, x[None, sl], [z[None, sl] for z in zs],
                                             self.states, want='target', targets=targets[None, sl])
            out[sl] = o[0]
            self.device_calls += 1
        return [1.0] + list(out[:T - 1])

    # ------------------------------------------------------------------ rate2
    def rate2(self, text, context=None):
        """rating.py:531-576: reset, then one (1,1) step per character."""
        if not context:
            context = self.underspecify_contexts()
        text = unicodedata.normalize('NFC', text)
        self.reset_states()
        entropy = 0
        result = []
        prev = 0
        for i, ch in enumerate(text):
            if ch not in self.mapping[0]:
                self.logger.error('unmapped character "%s" at input position %d', ch, i)
            idx = self._index(ch)
            if i == 0:
                result.append((ch, 1.0))
            else:
                probs, self.states = ln.step(self.w, np.array([prev]), [np.array([c]) for c in context],
                                             self.states)
                self.device_calls += 1
                prob = float(probs[0, idx])
                entropy -= math.log(max(prob, 1e-99), 2)
                result.append((ch, prob))
            prev = idx
        return result, pow(2.0, entropy / len(text))

    # ------------------------------------------------------------------ predict
    def predict(self, candidates, initial_states, context=None):
        """rating.py:578-639: one step for n hypotheses, explicit states; None -> zeros (in place)."""
        assert len(candidates) == len(initial_states)
        if not context:
            context = self.underspecify_contexts()
        n = len(candidates)
        x = np.array([self._index(c) for c in candidates], dtype=np.int64)
        zs = [np.full(n, c, dtype=np.int64) for c in context]
        for i, st in enumerate(initial_states):
            if not st:
                initial_states[i] = [np.zeros(self.width, dtype=float) for _ in range(2 * self.depth)]
        stacked = [np.vstack([st[k] for st in initial_states]).astype(self.w.E.dtype)
                   for k in range(2 * self.depth)]
        probs, new_states = ln.step(self.w, x, zs, stacked)
        self.device_calls += 1
        return ([probs[i, :] for i in range(n)],
                [[s[i:i + 1] for s in new_states] for i in range(n)])

    # ------------------------------------------------------------------ generate
    def generate(self, prefix, length, context=None, variants=1):
        """rating.py:642-709."""
        if not context:
            context = self.underspecify_contexts()
        states = [None]
        for ch in prefix[:-1]:
            _, states = self.predict([ch], states, context=context)
        frontier = [OracleNode(state=states[0], value=prefix[-1], cost=0.0)]
        for _ in range(length):
            current = frontier
            preds, new_states = self.predict([n.value for n in current], [n.state for n in current],
                                             context=context)
            frontier = []
            for j, node in enumerate(current):
                p = preds[j]
                best = np.argsort(p)[-10:]                                   # :695
                best = best[np.searchsorted(p[best], 0.004):]                # :696
                costs = -np.log(p[best])                                     # natural log (:697)
                for idx, cost in zip(best, costs):
                    if idx not in self.mapping[1]:                           # :700
                        continue
                    bisect.insort_left(frontier, OracleNode(parent=node, state=new_states[j],
                                                            value=self.mapping[1][idx], cost=cost))
            frontier = frontier[:256]                                        # :704
        return [''.join(n.value for n in res.to_sequence()) for res in frontier[0:variants]]

    # ------------------------------------------------------------------ rate_best
    @staticmethod
    def _edges_in_search_order(graph, start):
        """rating.py:763-770 (quirk: ``nodes`` is a growing list, duplicates allowed)."""
        import networkx as nx
        reached = [start]
        for tail in nx.topological_sort(graph):
            for head, _ in graph.in_edges([tail]):
                if head in reached:
                    yield head, tail
                    reached.append(tail)

    def rate_best(self, graph, start_node, end_node, start_traceback=None, context=None,
                  lm_weight=0.5, beam_width=10, beam_clustering_dist=0):
        """rating.py:712-859."""
        if not start_traceback:
            root = OracleNode(state=None, value='\n', cost=0.0)              # :760-762
            start_traceback = ([root], root)
        graph.nodes[start_node]['traceback'] = start_traceback[0]
        tail = 0
        tail_attrs = None
        for head, tail in self._edges_in_search_order(graph, start_node):
            edge = graph.edges[head, tail]
            element, alts = edge['element'], edge['alternatives']
            head_attrs, tail_attrs = graph.nodes[head], graph.nodes[tail]
            assert 'traceback' in head_attrs
            finished = tail_attrs['traceback'] if 'traceback' in tail_attrs else []
            pending = [OracleNode(parent=hyp, state=hyp.state, value="", cost=0.0, extras=(element, alt))
                       for hyp in head_attrs['traceback'] for alt in alts]   # :789-795
            complained = dict()
            rounds = 3 * max(len(alt.Unicode) for alt in alts)                # :797
            for _ in range(rounds):
                batch = []
                while pending:                                               # :800-810
                    cand = pending.pop()
                    if cand.value == cand.extras[1].Unicode:
                        if beam_clustering_dist and self._history_clustering(cand, finished,
                                                                             beam_clustering_dist):
                            continue
                        bisect.insort_left(finished, cand)
                    else:
                        bisect.insort_left(batch, cand)
                    if len(batch) >= self.batch_size:
                        break
                if not batch:
                    break
                if finished and batch[0].cum_cost >= finished[0].cum_cost + 15:   # :816
                    break
                preds, states = self.predict(
                    [c.value[-1] if c.value else c.parent.value[-1] for c in batch],
                    [c.state for c in batch], context)
                for i, cand in enumerate(batch):
                    alt = cand.extras[1]
                    ch = alt.Unicode[len(cand.value)]
                    if ch not in self.mapping[0]:
                        seen = complained.setdefault(alt.index, [])
                        if ch not in seen:
                            self.logger.error('unmapped character "%s" at input alternative %d of element %s',
                                              ch, alt.index or i, element.id if element else "space")
                            seen.append(ch)
                    idx = self._index(ch)
                    cand.cum_cost += (-math.log(max(preds[i][idx], 1e-99), 2) * lm_weight +
                                      -math.log(max(alt.conf, 1e-99), 2) * (1. - lm_weight))   # :839-840
                    cand.value += ch
                    cand.state = states[i]
                    if pending and cand.cum_cost >= pending[0].cum_cost + 2.5:    # :847
                        continue
                    bisect.insort_left(pending, cand)
                pending = pending[:rounds * self.batch_size]                      # :851
            tail_attrs['traceback'] = finished[:beam_width]                       # :852
        assert tail == end_node
        assert tail_attrs is not None and 'traceback' in tail_attrs
        return self.next_path(tail_attrs['traceback'], start_traceback)

    def next_path(self, beam, traceback):
        """rating.py:862-885."""
        prev_beam, prev_start = traceback
        # history of the best hypothesis from its (cut) root up to and including the first
        # ancestor that is a member of the previous traceback: the decided path of the
        # PREVIOUS graph; its last node is where all surviving hypotheses must descend from.
        trail = beam[0].to_sequence(stop_at=prev_beam)
        anchor = trail[-1]
        result = []
        for node in trail:
            if node.extras:
                element, alt = node.extras
                base = node.parent.cum_cost if node.parent else prev_start.cum_cost
                result.append((element, alt, pow(2.0, -(node.cum_cost - base) / len(alt.Unicode))))
        survivors = []
        for hyp in beam:
            if not hyp.to_sequence(stop_at=[anchor]):
                continue
            hyp.cut_at(anchor)
            bisect.insort_left(survivors, hyp)
        return result, anchor.cum_cost - prev_start.cum_cost, (survivors, anchor)

    def _history_clustering(self, cand, finished, distance=5):
        """rating.py:887-916 (compares only the first ``depth`` entries of the state list)."""
        for other in finished:
            if (cand.value == other.value and
                    all(np.linalg.norm(cand.state[k] - other.state[k]) < distance
                        for k in range(self.depth))):
                if other.cum_cost < cand.cum_cost:
                    return True
                finished.remove(other)
                break
        return False

    # ------------------------------------------------------------------ test
    def test_texts(self, items):
        """Perplexity as Rater.test computes it (rating.py:462-491) for [(text, context)] items:
        reset once; per window mean over all L positions of the cross-entropy (padding rows have an
        all-zero one-hot target and contribute 0); mean over windows; exp."""
        self.reset_states()
        losses = []
        L = self.length
        for text, context in items:
            text = unicodedata.normalize('NFC', text)
            T = len(text)
            if T <= 1:
                continue
            probs = self.rate(text, context)[1:]
            ce = -np.log(np.clip(np.asarray(probs, dtype=np.float64), 1e-7, 1 - 1e-7))
            n_windows = -(-(T - 1) // L)
            for k in range(n_windows):
                losses.append(ce[k * L:(k + 1) * L].sum() / L)
        return math.exp(float(np.mean(losses)))
