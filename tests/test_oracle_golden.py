"""The travelling CPU restatement (oracle/) against vectors produced by the reference's own
rating.py (tests/golden/make_golden.py).  Same float32 cell math underneath, so host logic
(windowing, padding, state carry, beam pruning, path selection) must agree exactly."""
import logging

import numpy as np
import pytest

from oracle import lstm_numpy as ln
from oracle.rater_oracle import OracleRater, OracleNode
from ocrd_keraslm_b200 import synth
from tests import util

quiet = logging.getLogger('oracle-test')
quiet.setLevel(logging.CRITICAL)


def make_oracle(length, batch_size=128):
    wd, maparr, mapping = util.small_model()
    return OracleRater(util.to_model(wd), mapping, length, batch_size=batch_size, logger=quiet)


def test_rate_windows_padding_and_state_carry():
    j, _ = util.golden()
    g = j['rate_small']
    o = make_oracle(g['length'])
    for case in g['cases']:            # successive calls, no reset: state + padded tail carry
        probs = o.rate(case['text'], case['context'])
        assert len(probs) == len(case['text'])
        np.testing.assert_allclose(np.array(probs, dtype=np.float64), np.array(case['probs']), rtol=1e-6, atol=1e-9)


def test_rate_degenerate_lengths():
    o = make_oracle(8)
    assert o.rate('') == []
    assert o.rate('a') == [1.0]


def test_rate2():
    j, _ = util.golden()
    g = j['rate2_small']
    o = make_oracle(g['length'])
    for case in g['cases']:
        res, ppl = o.rate2(case['text'], case['context'])
        assert [c for c, _ in res] == [c for c, _ in case['result']]
        np.testing.assert_allclose([p for _, p in res], [p for _, p in case['result']], rtol=1e-6)
        assert ppl == pytest.approx(case['perplexity'], rel=1e-6)


def test_rate_equals_rate2_from_zero_state():
    o = make_oracle(8)
    text = u'der ſchöne tag Z war'
    o.reset_states()
    a = o.rate(text, [7])
    b, _ = o.rate2(text, [7])
    np.testing.assert_allclose(np.array(a, dtype=np.float64), [p for _, p in b], rtol=1e-6)


def test_predict():
    j, a = util.golden()
    g = j['predict_small']
    o = make_oracle(8)
    s_in = [None] * 5
    p1, s1 = o.predict(list(g['cands1']), s_in, g['ctx1'])
    assert all(s is not None for s in s_in)          # mutated in place (rating.py:627)
    np.testing.assert_allclose(np.stack(p1), a['predict_small/probs1'], rtol=1e-6)
    np.testing.assert_allclose(np.stack([np.concatenate(s, 0) for s in s1]), a['predict_small/states1'], rtol=1e-6, atol=1e-8)
    st = [s1[k] if k >= 0 else None for k in g['perm2']]
    p2, s2 = o.predict(list(g['cands2']), st, g['ctx2'])
    np.testing.assert_allclose(np.stack(p2), a['predict_small/probs2'], rtol=1e-6)
    np.testing.assert_allclose(np.stack([np.concatenate(s, 0) for s in s2]), a['predict_small/states2'], rtol=1e-6, atol=1e-8)
    assert s2[0][0].shape == (1, 16)


def test_generate():
    j, _ = util.golden()
    g = j['generate_small']
    o = make_oracle(8)
    assert o.generate(g['prefix'], g['length'], context=g['context'], variants=g['variants']) == g['result']


def test_test_perplexity():
    j, _ = util.golden()
    g = j['test_small']
    o = make_oracle(g['length'])
    import math, os
    items = []
    for name, content in g['files']:
        parts = os.path.basename(name).split('.')[0].split('_')
        year = math.ceil(int(parts[2]) / 10) if len(parts) == 3 else 0
        items.append((content, [year]))
    assert o.test_texts(items) == pytest.approx(g['perplexity'], rel=1e-5)


@pytest.mark.parametrize('idx', range(4))
def test_rate_best_paths(idx):
    j, _ = util.golden()
    g = j['rate_best_small'][idx]
    o = make_oracle(8, batch_size=g['batch_size'])
    tb = None
    for page in g['pages']:
        graph = util.lattice_from_json(page['edges'])
        calls0 = o.device_calls
        path, entropy, tb = o.rate_best(graph, page['start'], page['end'], start_traceback=tb,
                                        context=g['context'], lm_weight=g['lm_weight'],
                                        beam_width=g['beam_width'], beam_clustering_dist=g['dist'])
        got = util.path_to_json(path)
        assert [p[:3] for p in got] == [p[:3] for p in page['path']]
        np.testing.assert_allclose([p[3] for p in got], [p[3] for p in page['path']], rtol=1e-9)
        assert entropy == pytest.approx(page['entropy'], rel=1e-12, abs=1e-12)
        assert [[n.cum_cost, n.value, n.length] for n in tb[0]] == [[c, v, l] for c, v, l in page['traceback']]
        assert tb[1].cum_cost == pytest.approx(page['start_cost'], rel=1e-12)
        assert o.device_calls - calls0 == page['predict_calls']
    path, entropy, _ = o.next_path(tb[0], ([], tb[1]))
    got = util.path_to_json(path)
    assert [p[:3] for p in got] == [p[:3] for p in g['flush']['path']]
    assert entropy == pytest.approx(g['flush']['entropy'], rel=1e-12)


def test_node_semantics():
    a = OracleNode(None, 'x', 1.0)
    b = OracleNode(None, 'y', 1.0)
    assert a == b and a in [b]
    with pytest.raises(TypeError):
        hash(a)
    c = OracleNode(None, 'z', 0.5, parent=a, extras=(None, synth.Alt('zzz', 0.9)))
    assert c.cum_cost == 1.5 and c.length == 2 and c.pro_cost() == 1.5 + 0.5 * 2
    assert c.to_sequence() == [a, c]


def test_cfg1_seeded_model_against_reference_run():
    j, a = util.golden()
    g = j['cfg1']
    wd = synth.weights(g['seed_weights'], width=128, depth=2, voc_size=256, style='trained')
    assert util.weights_checksum(wd) == pytest.approx(float(a['cfg1/weights_checksum'][0]), rel=1e-6)
    o = OracleRater(util.to_model(wd), synth.mapping_from_array(synth.vocabulary()), g['length'], logger=quiet)
    text = synth.text(g['seed_text'], g['size'])
    pa = o.rate(text[:g['split']], g['context'])
    pb = o.rate(text[g['split']:], g['context'])
    np.testing.assert_allclose(np.array(pa, dtype=np.float32), a['cfg1/probs_a'], rtol=1e-5)
    np.testing.assert_allclose(np.array(pb, dtype=np.float32), a['cfg1/probs_b'], rtol=1e-5)


def test_cfg4_three_contexts():
    j, a = util.golden()
    g = j['cfg4']
    wd = synth.weights(g['seed_weights'], width=32, depth=3, voc_size=256, contexts=((200, 16),) * 3, style='trained')
    assert util.weights_checksum(wd) == pytest.approx(float(a['cfg4/weights_checksum'][0]), rel=1e-6)
    o = OracleRater(util.to_model(wd), synth.mapping_from_array(synth.vocabulary()), g['length'], logger=quiet)
    for d, doc in enumerate(g['docs']):
        o.reset_states()
        probs = o.rate(synth.text(doc['seed_text'], doc['size']), doc['context'])
        np.testing.assert_allclose(np.array(probs, dtype=np.float32), a['cfg4/probs_%d' % d], rtol=1e-5)


def test_cell_math_against_torch_lstm():
    """Independent pin of the restated Keras cell arithmetic: torch.nn.LSTM (gate order i,f,g,o)."""
    import torch
    wd = synth.weights(5, width=24, depth=2, voc_size=50, contexts=((200, 10),), style='trained')
    m = util.to_model(wd)
    rng = np.random.default_rng(0)
    n, T = 3, 20
    x = rng.integers(0, 50, (n, T))
    z = rng.integers(0, 200, (n, T))
    probs, states = ln.run_sequence(m, x, [z], ln.zero_states(m, n), want='probs')
    lstm = torch.nn.LSTM(24 + 10, 24, num_layers=1, batch_first=True)
    inp = torch.from_numpy(ln.embed(m, x, [z]))
    hs = []
    for l in range(2):
        lstm = torch.nn.LSTM(inp.shape[-1], 24, num_layers=1, batch_first=True)
        with torch.no_grad():
            lstm.weight_ih_l0.copy_(torch.from_numpy(m.K[l].T))
            lstm.weight_hh_l0.copy_(torch.from_numpy(m.R[l].T))
            lstm.bias_ih_l0.copy_(torch.from_numpy(m.b[l]))
            lstm.bias_hh_l0.zero_()
            inp, (h, c) = lstm(inp)
        np.testing.assert_allclose(h[0].numpy(), states[2 * l], rtol=2e-5, atol=2e-6)
        np.testing.assert_allclose(c[0].numpy(), states[2 * l + 1], rtol=2e-5, atol=2e-6)
    ref = torch.softmax(inp @ torch.from_numpy(m.E).T, dim=-1).numpy()
    np.testing.assert_allclose(probs, ref, rtol=1e-4, atol=1e-7)


def test_torch_cpu_baseline_matches_numpy_oracle():
    """bench.py's second CPU line (oracle/lstm_torch_cpu.py: all streams batched, torch CPU) computes what the
    numpy oracle computes: log-probabilities of the targets and the final states, 3 context variables."""
    from oracle.lstm_torch_cpu import TorchCpuModel
    wd = synth.weights(11, width=40, depth=3, voc_size=64, contexts=((200, 16),) * 3, style='trained')
    m = util.to_model(wd)
    rng = np.random.default_rng(3)
    n, T = 5, 33
    x = rng.integers(0, 64, (n, T))
    y = rng.integers(0, 64, (n, T))
    z = [rng.integers(0, 200, n) for _ in range(3)]
    p, states = ln.run_sequence(m.astype(np.float64), x, [np.repeat(v[:, None], T, axis=1) for v in z],
                                ln.zero_states(m, n, np.float64), want='target', targets=y)
    logp, st = TorchCpuModel(m).rate(x, z, y)
    np.testing.assert_allclose(logp, np.log(p), atol=1e-4)          # the bar of the GPU path is 1e-3 nats
    for a, b in zip(st, states):
        np.testing.assert_allclose(a.numpy(), b, rtol=1e-4, atol=1e-5)
