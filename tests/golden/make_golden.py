#!/usr/bin/env python
"""Generate the golden vectors under tests/golden/ by running the REFERENCE's own host logic.

Runs only in the build container (needs /root/reference).  The reference module
/root/reference/ocrd_keraslm/lib/rating.py is imported in place (h5py stubbed, nothing
copied) and its ``Rater.rate / rate2 / predict / rate_best / next_path / generate / test``
are executed unmodified on top of ``oracle.keras_shim.KerasModelShim`` (float32 numpy
restatement of the Keras graph).  Inputs, weights and outputs are stored so that

  * ``oracle.rater_oracle`` (the travelling CPU restatement) and
  * the CUDA product path

can be checked against them on any box, without /root/reference.

    python tests/golden/make_golden.py        # rewrites tests/golden/*.npz, *.json
"""
import io
import json
import logging
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import keras_shim, lstm_numpy as ln       # noqa: E402
from ocrd_keraslm_b200 import synth                    # noqa: E402

SMALL_CHARS = u"\n abcdefghijklmnopqrstuvwxyzäöüſß.,-ABCDEFGH"


def small_mapping():
    cps = sorted(set(ord(c) for c in SMALL_CHARS))
    return np.array([0] + cps, dtype=np.uint32)


def to_model(wd):
    nctx = sum(1 for k in wd if k.startswith('context'))
    depth = sum(1 for k in wd if k.startswith('lstm_'))
    return ln.ModelWeights(wd['char_embedding'],
                           [wd['context%d_embedding' % (k + 1)] for k in range(nctx)],
                           [wd['lstm_%d' % (l + 1)][0] for l in range(depth)],
                           [wd['lstm_%d' % (l + 1)][1] for l in range(depth)],
                           [wd['lstm_%d' % (l + 1)][2] for l in range(depth)])


def flatten_weights(wd):
    flat = {}
    for k, v in wd.items():
        if isinstance(v, tuple):
            flat[k + '/kernel'], flat[k + '/recurrent_kernel'], flat[k + '/bias'] = v
        else:
            flat[k + '/embeddings'] = v
    return flat


def small_text(rng, n, unmapped=0.03):
    alphabet = list(SMALL_CHARS[2:])
    out = []
    for _ in range(n):
        r = rng.random()
        if r < 0.15:
            out.append(' ')
        elif r < 0.17:
            out.append('\n')
        elif r < 0.17 + unmapped:
            out.append('Z')         # not in SMALL_CHARS
        else:
            out.append(alphabet[int(rng.integers(0, len(alphabet)))])
    return ''.join(out)


def small_lattice(rng, n_words, n_alts, tag):
    import networkx as nx
    letters = list(u"abcdefghijklmnopqrstuvwxyzäöüſß")
    g = nx.DiGraph()
    g.add_node(0)
    node = 0
    for wi in range(n_words):
        n = int(rng.integers(1, 7))
        base = ''.join(letters[int(i)] for i in rng.integers(0, len(letters), n))
        alts = [synth.Alt(base, float(np.round(0.55 + 0.4 * rng.random(), 2)), 0)]
        for a in range(1, int(rng.integers(1, n_alts + 1))):
            kind = rng.random()
            pos = int(rng.integers(0, n))
            sub = letters[int(rng.integers(0, len(letters)))]
            if kind < 0.6:
                var = base[:pos] + sub + base[pos + 1:]
            elif kind < 0.8:
                var = base[:pos] + sub + base[pos:]           # insertion (longer alternative)
            elif kind < 0.9 and n > 1:
                var = base[:pos] + base[pos + 1:]             # deletion
            else:
                var = base[:pos] + 'Z' + base[pos + 1:]       # unmapped char
            alts.append(synth.Alt(var, float(np.round(0.3 + 0.6 * rng.random(), 2)), a))
        g.add_edge(node, node + 1, element=synth.Element('%s_w%03d' % (tag, wi)), alternatives=alts)
        node += 1
        sep = '\n' if (wi + 1) % 5 == 0 else ' '
        g.add_edge(node, node + 1, element=None, alternatives=[synth.Alt(sep, 1.0, 0)])
        node += 1
    return g, 0, node


def lattice_to_json(g):
    edges = []
    for u, v, d in g.edges(data=True):
        edges.append({'u': u, 'v': v, 'element': d['element'].id if d['element'] else None,
                      'alternatives': [[a.Unicode, a.conf, a.index] for a in d['alternatives']]})
    return edges


def path_to_json(path):
    return [[el.id if el else None, te.Unicode, te.index, float(score)] for el, te, score in path]


def beam_to_json(beam):
    return [[float(n.cum_cost), n.value, n.length] for n in beam]


class NamedFile(io.StringIO):
    def __init__(self, name, content):
        super().__init__(content)
        self.name = name


def main():
    logging.basicConfig(level=logging.CRITICAL)
    ref = keras_shim.import_reference_rating()
    out_json = {}
    arrays = {}

    # ------------------------------------------------------------------ small model (weights stored)
    maparr = small_mapping()
    mapping = synth.mapping_from_array(maparr)
    V = len(maparr)
    wd = synth.weights(1234, width=16, depth=2, voc_size=V, contexts=((200, 10),), style='trained')
    for k, v in flatten_weights(wd).items():
        arrays['small/' + k] = v
    arrays['small/mapping'] = maparr
    model = to_model(wd)
    L = 8
    rng = np.random.default_rng(99)

    # rate: successive calls WITHOUT reset (state + padded tail carry over), ragged lengths
    rater = keras_shim.make_reference_rater(ref, model, mapping, L, stateful=True, incremental=False)
    cases = []
    for T, ctx in [(2, [7]), (5, [7]), (8, None), (9, [0]), (10, [179]), (11, [7]), (30, [12]), (17, None), (64, [199])]:
        text = small_text(rng, T)
        probs = rater.rate(text, ctx)
        cases.append({'text': text, 'context': ctx, 'probs': [float(p) for p in probs]})
    out_json['rate_small'] = {'length': L, 'cases': cases}

    # rate2 (resets state itself)
    cases = []
    for T, ctx in [(1, [3]), (2, None), (13, [7]), (40, [150])]:
        text = small_text(rng, T)
        res, ppl = rater.rate2(text, ctx)
        cases.append({'text': text, 'context': ctx, 'result': [[c, float(p)] for c, p in res], 'perplexity': float(ppl)})
    out_json['rate2_small'] = {'length': L, 'cases': cases}

    # test (perplexity over "files", year context from file name)
    files = [NamedFile('/tmp/kant_aufklaerung_1784.txt', small_text(rng, 50)),
             NamedFile('/tmp/noyear.txt', small_text(rng, 21)),
             NamedFile('/tmp/goethe_faust_1808.txt', small_text(rng, 33))]
    ppl = rater.test(files)
    out_json['test_small'] = {'length': L, 'perplexity': float(ppl),
                              'files': [[f.name, f.getvalue()] for f in files]}

    # predict: None states, then feed the returned states back with other candidates
    inc = keras_shim.make_reference_rater(ref, model, mapping, L, stateful=False, incremental=True)
    cands1 = ['a', 'Z', '\n', 'ſ', 'b']
    probs1, states1 = inc.predict(list(cands1), [None] * 5, [7])
    cands2 = ['c', 'd', 'e', ' ', 'Z']
    st_in = [states1[4], states1[0], None, states1[2], states1[1]]
    probs2, states2 = inc.predict(list(cands2), list(st_in), None)
    arrays['predict_small/probs1'] = np.stack(probs1)
    arrays['predict_small/states1'] = np.stack([np.concatenate(s, axis=0) for s in states1])  # [n,2D,W]
    arrays['predict_small/probs2'] = np.stack(probs2)
    arrays['predict_small/states2'] = np.stack([np.concatenate(s, axis=0) for s in states2])
    out_json['predict_small'] = {'cands1': cands1, 'ctx1': [7], 'cands2': cands2, 'ctx2': None,
                                 'perm2': [4, 0, -1, 2, 1]}

    # generate
    gens = inc.generate(u'der ', 12, context=[17], variants=3)
    out_json['generate_small'] = {'prefix': u'der ', 'length': 12, 'context': [17], 'variants': 3, 'result': gens}

    # rate_best: chained pages, several settings
    rb = []
    for name, beam_width, dist, lmw, n_words, n_alts, n_pages, bs in [
            ('bw5', 5, 0, 0.5, 10, 3, 3, 128),
            ('bw10_cluster', 10, 5, 0.5, 10, 4, 3, 128),
            ('bw100_wide', 100, 0, 0.7, 8, 5, 2, 128),
            ('bw20_smallbatch', 20, 0, 0.5, 8, 4, 2, 16)]:
        inc = keras_shim.make_reference_rater(ref, model, mapping, L, stateful=False, incremental=True)
        inc.batch_size = bs
        lrng = np.random.default_rng(abs(hash(name)) % 1000 if False else sum(map(ord, name)))
        tb = None
        pages = []
        for pg in range(n_pages):
            g, s, e = small_lattice(lrng, n_words, n_alts, 'p%d' % pg)
            edges = lattice_to_json(g)
            calls0 = inc.model.calls
            path, entropy, tb = inc.rate_best(g, s, e, start_traceback=tb, context=[17],
                                              lm_weight=lmw, beam_width=beam_width, beam_clustering_dist=dist)
            pages.append({'edges': edges, 'start': s, 'end': e, 'path': path_to_json(path),
                          'entropy': float(entropy), 'traceback': beam_to_json(tb[0]),
                          'start_cost': float(tb[1].cum_cost), 'predict_calls': inc.model.calls - calls0})
        path, entropy, tb2 = inc.next_path(tb[0], ([], tb[1]))   # flush last page (wrapper/rate.py:222)
        rb.append({'name': name, 'beam_width': beam_width, 'dist': dist, 'lm_weight': lmw,
                   'batch_size': bs, 'context': [17], 'pages': pages,
                   'flush': {'path': path_to_json(path), 'entropy': float(entropy)}})
    out_json['rate_best_small'] = rb

    # ------------------------------------------------------------------ config-1 sized model (seeded, not stored)
    maparr = synth.vocabulary()
    mapping = synth.mapping_from_array(maparr)
    wd = synth.weights(7 + 1, width=128, depth=2, voc_size=256, style='trained')
    model = to_model(wd)
    arrays['cfg1/weights_checksum'] = np.array([float(np.sum([np.abs(v).sum() for v in flatten_weights(wd).values()]))])
    rater = keras_shim.make_reference_rater(ref, model, mapping, 256, stateful=True, incremental=False)
    text = synth.text(20261104 + 1000, 1500)
    probs_a = rater.rate(text[:700], [179])
    probs_b = rater.rate(text[700:], [179])       # continues from carried state incl. padded tail
    arrays['cfg1/probs_a'] = np.array(probs_a, dtype=np.float32)
    arrays['cfg1/probs_b'] = np.array(probs_b, dtype=np.float32)
    out_json['cfg1'] = {'seed_weights': 8, 'seed_text': 20261104 + 1000, 'size': 1500, 'split': 700,
                        'context': [179], 'length': 256}

    # context-conditioned model (config-4 shaped, small): 3 context tables of dim 16
    wd = synth.weights(7 + 4, width=32, depth=3, voc_size=256, contexts=((200, 16),) * 3, style='trained')
    model = to_model(wd)
    arrays['cfg4/weights_checksum'] = np.array([float(np.sum([np.abs(v).sum() for v in flatten_weights(wd).values()]))])
    rater = keras_shim.make_reference_rater(ref, model, mapping, 64, stateful=True, incremental=False)
    docs = []
    for d, ctx in enumerate([[3, 0, 199], [0, 0, 0], [17, 42, 1]]):
        rater.model.reset_states()
        text = synth.text(20261104 + 4000 + d, 150 + 37 * d)
        probs = rater.rate(text, ctx)
        arrays['cfg4/probs_%d' % d] = np.array(probs, dtype=np.float32)
        docs.append({'seed_text': 20261104 + 4000 + d, 'size': 150 + 37 * d, 'context': ctx})
    out_json['cfg4'] = {'seed_weights': 11, 'length': 64, 'docs': docs}

    np.savez_compressed(os.path.join(HERE, 'golden.npz'), **arrays)
    with open(os.path.join(HERE, 'golden.json'), 'w', encoding='utf-8') as f:
        json.dump(out_json, f, ensure_ascii=False, indent=1)
    print('wrote', os.path.join(HERE, 'golden.npz'), os.path.join(HERE, 'golden.json'))


if __name__ == '__main__':
    main()
