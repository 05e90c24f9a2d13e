"""Shared helpers for the test-suite: golden loading, model construction, lattices."""
import json
import os

import numpy as np

from oracle import lstm_numpy as ln
from ocrd_keraslm_b200 import synth

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')

_cache = {}


def golden():
    if 'g' not in _cache:
        with open(os.path.join(GOLDEN_DIR, 'golden.json'), encoding='utf-8') as f:
            j = json.load(f)
        a = dict(np.load(os.path.join(GOLDEN_DIR, 'golden.npz')))
        _cache['g'] = (j, a)
    return _cache['g']


def weights_dict_from_flat(arrays, prefix):
    """{'char_embedding': E, 'context1_embedding': C, 'lstm_1': (K,R,b), ...} from the npz."""
    out = {}
    groups = {}
    for k, v in arrays.items():
        if not k.startswith(prefix + '/'):
            continue
        parts = k[len(prefix) + 1:].split('/')
        if len(parts) != 2:
            continue
        groups.setdefault(parts[0], {})[parts[1]] = v
    for layer, d in groups.items():
        if 'embeddings' in d:
            out[layer] = d['embeddings']
        else:
            out[layer] = (d['kernel'], d['recurrent_kernel'], d['bias'])
    return out


def to_model(wd, dtype=None):
    nctx = sum(1 for k in wd if k.startswith('context'))
    depth = sum(1 for k in wd if k.startswith('lstm_'))
    m = ln.ModelWeights(wd['char_embedding'],
                        [wd['context%d_embedding' % (k + 1)] for k in range(nctx)],
                        [wd['lstm_%d' % (l + 1)][0] for l in range(depth)],
                        [wd['lstm_%d' % (l + 1)][1] for l in range(depth)],
                        [wd['lstm_%d' % (l + 1)][2] for l in range(depth)])
    return m.astype(dtype) if dtype is not None else m


def small_model():
    j, a = golden()
    wd = weights_dict_from_flat(a, 'small')
    maparr = a['small/mapping']
    return wd, maparr, synth.mapping_from_array(maparr)


def weights_checksum(wd):
    tot = 0.0
    for v in wd.values():
        for arr in (v if isinstance(v, tuple) else (v,)):
            tot += float(np.abs(arr).sum())
    return tot


def lattice_from_json(edges):
    import networkx as nx
    g = nx.DiGraph()
    g.add_node(0)
    for e in edges:
        el = synth.Element(e['element']) if e['element'] is not None else None
        alts = [synth.Alt(t, c, i) for t, c, i in e['alternatives']]
        g.add_edge(e['u'], e['v'], element=el, alternatives=alts)
    return g


def path_to_json(path):
    return [[el.id if el else None, te.Unicode, te.index, float(score)] for el, te, score in path]
