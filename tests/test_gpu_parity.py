"""Parity of the CUDA path (through the C ABI, via the product ``Rater``) against

  * the golden vectors produced by the reference's own rating.py (tests/golden/), and
  * the CPU oracle on the same seeded inputs.

Tolerances (BASELINE.json north_star): per-character log-probabilities within 1e-3 nats,
hidden/cell states within 1e-4 relative, identical beam-search paths.
"""
import io
import logging
import math

import numpy as np
import pytest

from oracle import lstm_numpy as ln
from oracle.rater_oracle import OracleRater
from ocrd_keraslm_b200 import synth
from tests import util

pytestmark = pytest.mark.gpu

LOGP_TOL = 1e-3       # nats
STATE_TOL = 1e-4      # relative (max |diff| / max |ref|)

quiet = logging.getLogger('gpu-test')
quiet.setLevel(logging.CRITICAL)


def product(wd, maparr, length, stateful=True, incremental=False):
    from ocrd_keraslm_b200.lib import Rater
    return Rater.from_weights(wd, maparr, length, stateful=stateful, incremental=incremental, logger=quiet)


def assert_logp_close(got, ref, tol=LOGP_TOL):
    got = np.log(np.maximum(np.asarray(got, dtype=np.float64), 1e-300))
    ref = np.log(np.maximum(np.asarray(ref, dtype=np.float64), 1e-300))
    assert got.shape == ref.shape
    assert np.abs(got - ref).max() <= tol, np.abs(got - ref).max()


def assert_state_close(got, ref, tol=STATE_TOL):
    got, ref = np.asarray(got, np.float64), np.asarray(ref, np.float64)
    assert got.shape == ref.shape
    assert np.abs(got - ref).max() <= tol * max(np.abs(ref).max(), 1e-30), np.abs(got - ref).max()


# ------------------------------------------------------------------ golden vectors (small model)
def test_rate_golden_windows_padding_state_carry():
    j, _ = util.golden()
    g = j['rate_small']
    wd, maparr, _ = util.small_model()
    r = product(wd, maparr, g['length'])
    for case in g['cases']:                 # successive calls, no reset in between
        probs = r.rate(case['text'], case['context'])
        assert len(probs) == len(case['text']) and probs[0] == 1.0
        assert_logp_close(probs, case['probs'])


def test_rate_degenerate_lengths():
    wd, maparr, _ = util.small_model()
    r = product(wd, maparr, 8)
    assert r.rate('') == []
    assert r.rate('a') == [1.0]


def test_rate2_golden():
    j, _ = util.golden()
    g = j['rate2_small']
    wd, maparr, _ = util.small_model()
    r = product(wd, maparr, g['length'])
    for case in g['cases']:
        res, ppl = r.rate2(case['text'], case['context'])
        assert [c for c, _ in res] == [c for c, _ in case['result']]
        assert_logp_close([p for _, p in res], [p for _, p in case['result']])
        assert ppl == pytest.approx(case['perplexity'], rel=2e-3)


def test_predict_golden_probs_and_states():
    j, a = util.golden()
    g = j['predict_small']
    wd, maparr, _ = util.small_model()
    r = product(wd, maparr, 8, stateful=False, incremental=True)
    s_in = [None] * 5
    p1, s1 = r.predict(list(g['cands1']), s_in, g['ctx1'])
    assert all(s is not None for s in s_in)             # mutated in place like the reference
    assert_logp_close(np.stack(p1), a['predict_small/probs1'])
    assert_state_close(np.stack([np.concatenate(list(s), 0) for s in s1]), a['predict_small/states1'])
    st = [s1[k] if k >= 0 else None for k in g['perm2']]
    p2, s2 = r.predict(list(g['cands2']), st, g['ctx2'])
    assert_logp_close(np.stack(p2), a['predict_small/probs2'])
    assert_state_close(np.stack([np.concatenate(list(s), 0) for s in s2]), a['predict_small/states2'])
    assert len(s2[0]) == 4 and s2[0][0].shape == (1, 16)
    np.testing.assert_allclose(np.stack(p2).sum(axis=1), 1.0, rtol=1e-5)


def test_predict_accepts_host_state_lists():
    wd, maparr, _ = util.small_model()
    r = product(wd, maparr, 8, stateful=False, incremental=True)
    p1, s1 = r.predict(['a', 'b'], [None, None], [3])
    as_lists = [[np.array(x) for x in s] for s in s1]
    pa, _ = r.predict(['c', 'd'], list(s1), [3])
    pb, _ = r.predict(['c', 'd'], as_lists, [3])
    np.testing.assert_array_equal(np.stack(pa), np.stack(pb))


def test_generate_golden():
    j, _ = util.golden()
    g = j['generate_small']
    wd, maparr, _ = util.small_model()
    r = product(wd, maparr, 8, stateful=False, incremental=True)
    assert r.generate(g['prefix'], g['length'], context=g['context'], variants=g['variants']) == g['result']


def test_test_perplexity_golden():
    j, _ = util.golden()
    g = j['test_small']
    wd, maparr, _ = util.small_model()
    r = product(wd, maparr, g['length'])
    files = []
    for name, content in g['files']:
        f = io.StringIO(content)
        f.name = name
        files.append(f)
    assert r.test(files) == pytest.approx(g['perplexity'], rel=1e-3)


@pytest.mark.parametrize('idx', range(4))
def test_rate_best_golden_paths(idx):
    j, _ = util.golden()
    g = j['rate_best_small'][idx]
    wd, maparr, _ = util.small_model()
    r = product(wd, maparr, 8, stateful=False, incremental=True)
    r.batch_size = g['batch_size']
    tb = None
    for page in g['pages']:
        graph = util.lattice_from_json(page['edges'])
        path, entropy, tb = r.rate_best(graph, page['start'], page['end'], start_traceback=tb, context=g['context'],
                                        lm_weight=g['lm_weight'], beam_width=g['beam_width'],
                                        beam_clustering_dist=g['dist'])
        got = util.path_to_json(path)
        assert [p[:3] for p in got] == [p[:3] for p in page['path']]
        np.testing.assert_allclose([p[3] for p in got], [p[3] for p in page['path']], rtol=2e-3)
        assert entropy == pytest.approx(page['entropy'], rel=2e-3, abs=2e-3)
        assert [n.value for n in tb[0]] == [v for _, v, _ in page['traceback']]
        np.testing.assert_allclose([n.cum_cost for n in tb[0]], [c for c, _, _ in page['traceback']], rtol=2e-3, atol=2e-3)
    path, entropy, _ = r.next_path(tb[0], ([], tb[1]))
    assert [p[:3] for p in util.path_to_json(path)] == [p[:3] for p in g['flush']['path']]


# ------------------------------------------------------------------ seeded models vs reference-run goldens
def test_cfg1_model_against_reference_run():
    j, a = util.golden()
    g = j['cfg1']
    wd = synth.weights(g['seed_weights'], width=128, depth=2, voc_size=256, style='trained')
    r = product(wd, synth.vocabulary(), g['length'])
    text = synth.text(g['seed_text'], g['size'])
    pa = r.rate(text[:g['split']], g['context'])
    pb = r.rate(text[g['split']:], g['context'])
    assert_logp_close(pa, a['cfg1/probs_a'])
    assert_logp_close(pb, a['cfg1/probs_b'])


def test_cfg4_three_context_variables():
    j, a = util.golden()
    g = j['cfg4']
    wd = synth.weights(g['seed_weights'], width=32, depth=3, voc_size=256, contexts=((200, 16),) * 3, style='trained')
    r = product(wd, synth.vocabulary(), g['length'])
    for d, doc in enumerate(g['docs']):
        r.reset_states()
        probs = r.rate(synth.text(doc['seed_text'], doc['size']), doc['context'])
        assert_logp_close(probs, a['cfg4/probs_%d' % d])
    # the same documents as one mixed-context batch
    texts = [synth.text(doc['seed_text'], doc['size']) for doc in g['docs']]
    many = r.rate_many(texts, [doc['context'] for doc in g['docs']])
    for d in range(len(texts)):
        ref = np.array(a['cfg4/probs_%d' % d])
        assert_logp_close(many[d], ref[:len(many[d])])


def test_context_out_of_range_raises():
    wd, maparr, _ = util.small_model()
    r = product(wd, maparr, 8)
    with pytest.raises(ValueError):
        r.rate('abc', [200])


# ------------------------------------------------------------------ against the oracle on the same seeded inputs
def oracle_for(wd, maparr, length):
    return OracleRater(util.to_model(wd), synth.mapping_from_array(maparr), length, logger=quiet)


@pytest.mark.parametrize('width,depth,n_docs', [(64, 2, 3), (64, 2, 20), (128, 3, 70), (512, 3, 64)])
def test_rate_many_matches_oracle(width, depth, n_docs):
    """K-parallel (<= 8 streams) and stream-parallel kernels, ragged lengths, incl. BASELINE config 2's shape."""
    wd = synth.weights(11 + width, width=width, depth=depth, voc_size=256, style='trained')
    maparr = synth.vocabulary()
    r = product(wd, maparr, 32)
    rng = np.random.default_rng(width + n_docs)
    sizes = rng.integers(2, 60 if width < 512 else 24, n_docs)
    sizes[0] = 1
    texts = [synth.text(1000 + i, int(s)) for i, s in enumerate(sizes)]
    ctxs = [[int(rng.integers(0, 200))] for _ in texts]
    got = r.rate_many(texts, ctxs)
    o = oracle_for(wd, maparr, 32)
    for i, (t, c) in enumerate(zip(texts, ctxs)):
        ref, _ = o.rate2(t, c)
        assert_logp_close(got[i], [p for _, p in ref])


@pytest.mark.parametrize('width,depth,n_docs,nseg', [(64, 2, 5, 7), (128, 3, 40, 16), (512, 3, 64, 3)])
def test_rate_many_in_time_segments_matches_one_launch_and_oracle(width, depth, n_docs, nseg):
    """The pipelined forms of rate_many (worker thread; short texts: segments = groups of whole streams; long texts:
    segments along time, state carried in the pool slots) on inputs small enough for the oracle: both must equal
    the one-launch form bit for bit, and the oracle in tolerance.  Ragged lengths include streams that contribute
    nothing to some time segments (length < number of segments)."""
    wd = synth.weights(5 + width, width=width, depth=depth, voc_size=256, style='trained')
    maparr = synth.vocabulary()
    r = product(wd, maparr, 32)
    rng = np.random.default_rng(width * 3 + n_docs)
    sizes = rng.integers(2, 50 if width < 512 else 20, n_docs)
    sizes[0], sizes[1] = 1, 3
    texts = [synth.text(2000 + i, int(s)) for i, s in enumerate(sizes)]
    ctxs = [[int(rng.integers(0, 200))] for _ in texts]
    one = r.rate_many(texts, ctxs, return_logp=True)
    r.MAX_SEGMENTS, r.MIN_SEGMENT_POSITIONS = nseg, 1       # instance overrides: force nseg segments ...
    r.LAUNCH_STREAMS = 2                                     # ... of whole streams, cut at multiples of 2 (short texts)
    seg = r.rate_many(texts, ctxs, return_logp=True)
    r.GATHER_MAX_MEAN_LENGTH = 0                             # the host path for few long texts (per-text slices)
    sliced = r.rate_many(texts, ctxs, return_logp=True)
    for a, b, c in zip(one, seg, sliced):
        # groups of a few streams may run on another kernel (<= 8 streams: FP32 FMA) than the one launch of all
        np.testing.assert_allclose(a, b, rtol=0, atol=2e-5)
        np.testing.assert_array_equal(a, c)                  # time segments: same kernel, same arithmetic
    o = oracle_for(wd, maparr, 32)
    for i in (0, 1, 2, n_docs - 1):
        ref, _ = o.rate2(texts[i], ctxs[i])
        assert_logp_close(np.exp(seg[i]), [p for _, p in ref])


@pytest.mark.parametrize('n_docs', [1, 2, 5, 8])
def test_value_tick_exchange_kernel_over_many_ring_turns(n_docs):
    """<= 8 streams run on the barrier-free {value, tick} exchange kernel: ragged lengths from 1 to several
    hundred ticks (the exchange ring holds 32), against the oracle, twice in a row (ring words re-armed)."""
    wd = synth.weights(91, width=128, depth=3, voc_size=256, style='trained')
    maparr = synth.vocabulary()
    r = product(wd, maparr, 32)
    o = oracle_for(wd, maparr, 32)
    sizes = [700, 2, 33, 65, 310, 32, 129, 500][:n_docs]
    texts = [synth.text(3000 + i, s) for i, s in enumerate(sizes)]
    ctxs = [[i * 7] for i in range(n_docs)]
    first = r.rate_many(texts, ctxs)
    again = r.rate_many(texts, ctxs)
    for i, (t, c) in enumerate(zip(texts, ctxs)):
        np.testing.assert_array_equal(first[i], again[i])
        ref, _ = o.rate2(t, c)
        assert_logp_close(first[i], [p for _, p in ref])


@pytest.mark.parametrize('depth,voc', [(2, 256), (1, 256), (3, 256), (2, 100), (2, 400)])
def test_register_resident_cluster_kernel_matches_oracle(depth, voc):
    """klm_seq_rg: ONE text, width 128 -- the model in the registers of one 8-CTA cluster, recurrence through distributed
    shared memory (the reference's default use: Rater.rate / rate2 of a string).  Texts of several lengths incl. one and
    two positions, state carried from call to call (padded tail included), twice in a row bit-identical."""
    wd = synth.weights(141 + depth, width=128, depth=depth, voc_size=voc, style='trained')
    maparr = synth.vocabulary()[:voc] if voc <= 256 else np.concatenate(
        [synth.vocabulary(), np.arange(0x400, 0x400 + voc - 256, dtype=np.uint32)])
    r = product(wd, maparr, 32)
    assert r._native.info()['cl_cluster'] in (8, 16)
    o = oracle_for(wd, maparr, 32)
    for size, ctx in ((150, [7]), (2, [0]), (3, [199]), (40, [33]), (333, [5])):
        text = synth.text(3300 + size, size, mapping_arr=maparr)
        launches = r._native.info()['cl_launches']
        got, ppl = r.rate2(text, ctx)
        assert r._native.info()['cl_launches'] == launches + 1
        again, _ = r.rate2(text, ctx)
        assert got == again
        ref, ppl_o = o.rate2(text, ctx)
        assert_logp_close([q for _, q in got], [q for _, q in ref])
        assert ppl == pytest.approx(ppl_o, rel=1e-3)
    # stateful rating with carry (Rater.rate: windows of `length`, padded tail), against the oracle, incl. the final state
    r.reset_states()
    o.reset_states()
    text = synth.text(3399, 150, mapping_arr=maparr)
    for piece in (text[:70], text[70:], text):
        assert_logp_close(r.rate(piece, [11]), o.rate(piece, [11]))
    assert_state_close(r._native.state_to_host([0])[0], np.concatenate([s for s in o.states], axis=0))
    # a single hypothesis through klm_step (predict) takes the same kernel: full softmax row
    ri = product(wd, maparr, 32, stateful=False, incremental=True)
    probs, states = ri.predict(['a'], [None], context=[3])
    pr, st = o.predict(['a'], [None], context=[3])
    assert_logp_close(probs[0], pr[0], tol=2e-3)


def test_rate_many_degenerate_inputs():
    """Empty list, empty strings and one-character texts next to ordinary ones; non-BMP characters (one code point,
    unmapped -> logged, index 0 like the reference's unknown characters)."""
    wd, maparr, _ = util.small_model()
    r = product(wd, maparr, 8)
    assert r.rate_many([]) == []
    texts = ['', 'a', 'abc', '', 'ab\U0001F600cd', 'x']
    got = r.rate_many(texts)
    assert [len(g) for g in got] == [len(t) for t in texts]
    assert all(g[0] == 1.0 for g in got if len(g))
    o = oracle_for(wd, maparr, 8)
    for i in (2, 4):
        ref, _ = o.rate2(texts[i])
        assert_logp_close(got[i], [p for _, p in ref])
    r.GATHER_MAX_MEAN_LENGTH = 0              # the per-text slicing path
    again = r.rate_many(texts)
    for a, b in zip(got, again):
        np.testing.assert_array_equal(a, b)


def test_rate_many_reports_unmapped_characters_like_the_reference(caplog):
    """Unmapped characters are rated as index 0 and reported one by one with their position in the text (reference
    :1116-1118) -- from both host paths of rate_many (one lookup over all texts / per-text slices), and not at all
    when the logger does not take errors; the list of many short texts carries its scores as ONE array, too."""
    import logging
    wd, maparr, _ = util.small_model()
    r = product(wd, maparr, 8)
    r.logger = logging.getLogger('test_unmapped')
    texts = ['ab\u2603cd', 'xyz', '', '\u2603', 'q\u2603\u2603']
    want = [('\u2603', 2), ('\u2603', 0), ('\u2603', 1), ('\u2603', 2)]
    for limit in (4096, 0):                                 # gather path, per-text path
        r.GATHER_MAX_MEAN_LENGTH = limit
        caplog.clear()
        with caplog.at_level(logging.ERROR, logger='test_unmapped'):
            got = r.rate_many(texts)
        seen = [(rec.args[0], rec.args[1]) for rec in caplog.records if 'unmapped character' in rec.msg]
        assert sorted(seen) == sorted(want), (limit, seen)
        flat = getattr(got, 'flat', None)
        if limit:
            assert flat is not None
            np.testing.assert_array_equal(flat, np.concatenate(got))
        caplog.clear()
        r.logger.setLevel(logging.CRITICAL)
        again = r.rate_many(texts)
        r.logger.setLevel(logging.NOTSET)
        assert not [rec for rec in caplog.records if 'unmapped character' in rec.msg]
        for a, b in zip(got, again):
            np.testing.assert_array_equal(a, b)


def test_states_after_rate_match_oracle():
    wd = synth.weights(3, width=128, depth=3, voc_size=256, style='trained')
    maparr = synth.vocabulary()
    r = product(wd, maparr, 16)
    o = oracle_for(wd, maparr, 16)
    text = synth.text(77, 100)               # (T-1) % L != 0: padded tail advances the state
    assert_logp_close(r.rate(text, [5]), o.rate(text, [5]))
    got = r._native.state_to_host([0])[0]
    ref = np.concatenate([s for s in o.states], axis=0)
    assert_state_close(got, ref)


def test_long_document_drift_and_chunking():
    """Size-independent properties at BASELINE config 1's size: rate == rate2 from a zero state, and rating
    a document in window-aligned pieces (state carry) equals rating it at once."""
    wd = synth.weights(8, width=128, depth=2, voc_size=256, style='trained')
    maparr = synth.vocabulary()
    L = 256
    r = product(wd, maparr, L)
    text = synth.text(20261104 + 1000, 100000)
    r.reset_states()
    whole = np.array(r.rate(text, [179]))
    two, ppl = r.rate2(text, [179])
    np.testing.assert_array_equal(whole, np.array([p for _, p in two], dtype=np.float32))
    r.reset_states()
    cut = 40 * L + 1                          # first piece consumes exactly 40 full windows
    a = r.rate(text[:cut], [179])
    b = r.rate(text[cut - 1:], [179])         # overlap by one char: b[0] is the dummy 1.0
    np.testing.assert_allclose(np.array(a + b[1:]), whole, rtol=1e-6)
    # oracle on a prefix (the oracle finishes 3000 chars in seconds)
    o = oracle_for(wd, maparr, L)
    ref = o.rate(text[:3000], [179])
    assert_logp_close(whole[:2999], ref[:2999])
    assert ppl > 1.0 and math.isfinite(ppl)


# ------------------------------------------------------------------ native beam edge (klm_beam_edge) vs the host-language loop
@pytest.mark.parametrize('width,beam_width,batch_size', [(64, 10, 128), (128, 100, 128), (128, 7, 16)])
def test_rate_best_native_edge_equals_python_loop(width, beam_width, batch_size):
    wd = synth.weights(21, width=width, depth=2, voc_size=256, style='trained')
    maparr = synth.vocabulary()
    results = []
    for native, lattice in ((True, True), (False, False), (True, False)):   # whole lattice natively / host loop / edge by edge
        r = product(wd, maparr, 8, stateful=False, incremental=True)
        r.native_beam, r.native_lattice = native, lattice
        r.batch_size = batch_size
        tb = None
        pages = []
        for pg in range(2):
            graph, start, end = synth.lattice(900 + pg, n_words=12, n_alts=5)
            path, entropy, tb = r.rate_best(graph, start, end, start_traceback=tb, context=[42], lm_weight=0.5,
                                            beam_width=beam_width, beam_clustering_dist=0)
            pages.append(([(el.id if el else None, alt.Unicode, alt.index) for el, alt, _ in path],
                          [s for _, _, s in path], entropy, [n.value for n in tb[0]], [n.cum_cost for n in tb[0]]))
        path, entropy, _ = r.next_path(tb[0], ([], tb[1]))
        pages.append(([(el.id if el else None, alt.Unicode, alt.index) for el, alt, _ in path], [s for _, _, s in path],
                      entropy, [], []))
        results.append((pages, r.device_steps, r.hypothesis_steps))
    (pa, sa, ha), (pb, sb, hb) = results[:2]
    for pc, sc, hc in results[2:]:
        assert (sc, hc) == (sb, hb)
        for x, y in zip(pc, pb):
            assert x[0] == y[0] and x[3] == y[3]
            np.testing.assert_allclose(x[4], y[4], rtol=1e-6, atol=1e-6)
    assert (sa, ha) == (sb, hb)                      # same number of device steps and hypotheses advanced
    for x, y in zip(pa, pb):
        assert x[0] == y[0] and x[3] == y[3]
        np.testing.assert_allclose(x[1], y[1], rtol=1e-6)
        np.testing.assert_allclose(x[4], y[4], rtol=1e-6, atol=1e-6)
        assert x[2] == pytest.approx(y[2], rel=1e-6, abs=1e-6)


# ------------------------------------------------------------------ vocabulary pruning
def test_remove_from_mapping_shifts_indices_and_rebuilds_model():
    """prune-charset (reference :416-460): the character's embedding row disappears, higher indices move down
    by one, and the rebuilt device model rates like the oracle on the pruned weights."""
    wd, maparr, mapping = util.small_model()
    r = product(wd, maparr, 8)
    victim = sorted(mapping[0].items(), key=lambda kv: kv[1])[3][0]       # some character in the middle
    vidx = mapping[0][victim]
    voc0 = r.voc_size
    assert r.remove_from_mapping(char='☃') is False                   # unmapped: logged, nothing changes
    assert r.remove_from_mapping() is False
    assert r.voc_size == voc0
    assert r.remove_from_mapping(char=victim) is True
    assert r.status == 2 and r.voc_size == voc0 - 1
    assert victim not in r.mapping[0]
    for c, i in mapping[0].items():
        if c == victim:
            continue
        assert r.mapping[0][c] == (i - 1 if i > vidx else i)
        assert r.mapping[1][r.mapping[0][c]] == c
    assert sorted(r.mapping[1]) == list(range(1, voc0 - 1))
    # oracle on the pruned weights and mapping
    wd2 = dict(wd)
    wd2['char_embedding'] = np.delete(np.asarray(wd['char_embedding']), vidx, 0)
    c_i = dict((c, (i - 1 if i > vidx else i)) for c, i in mapping[0].items() if c != victim)
    i_c = dict((i, c) for c, i in c_i.items())
    o = OracleRater(util.to_model(wd2), (c_i, i_c), 8, logger=quiet)
    text = ''.join(sorted(mapping[0], key=mapping[0].get)[:20]) * 3        # contains the victim: now unmapped (index 0)
    assert_logp_close(r.rate(text, [5]), o.rate(text, [5]))
    # removal by index
    assert r.remove_from_mapping(idx=1) is True
    assert r.voc_size == voc0 - 2


# ------------------------------------------------------------------ device top-k (klm_step_topk; Rater.generate)
@pytest.mark.parametrize('width,n,k', [(64, 5, 10), (128, 40, 10), (256, 130, 3), (64, 3, 256)])
def test_step_topk_matches_argsort_of_full_rows(width, n, k):
    """The k best entries of every softmax row, picked on the device, are what np.argsort(p)[-k:] (reference :690-692)
    picks from the full rows klm_step returns: same indices in the same (ascending) order, same probabilities; the new
    states are the same, too."""
    wd = synth.weights(33, width=width, depth=2, voc_size=256, style='trained')
    maparr = synth.vocabulary()
    r = product(wd, maparr, 8, stateful=False, incremental=True)
    nat = r._native
    rng = np.random.default_rng(n)
    chars = rng.integers(0, 256, n).astype(np.int32)
    ctx = rng.integers(0, 200, (n, 1)).astype(np.int32)
    s_in = np.full(n, -1, dtype=np.int32)
    a, b = r._pool.acquire(n), r._pool.acquire(n)
    full = np.empty((n, 256), dtype=np.float32)
    nat.step(chars, ctx, s_in, a, None, full)
    idx, prob = nat.step_topk(chars, ctx, s_in, b, k)
    for j in range(n):
        want = np.argsort(full[j], kind='stable')[-k:]
        np.testing.assert_array_equal(idx[j], want)
        np.testing.assert_array_equal(prob[j], full[j][want])
    np.testing.assert_array_equal(nat.state_to_host(a), nat.state_to_host(b))
