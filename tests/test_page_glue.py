"""PAGE-XML glue (reference wrapper/rate.py): graph construction, write-back and the consistency
of the higher levels, driven by the CPU oracle here and by the CUDA path against the oracle in the gpu test."""
import logging
import math

import numpy as np
import pytest

from oracle.rater_oracle import OracleRater
from ocrd_keraslm_b200 import synth
from ocrd_keraslm_b200.wrapper import page as pg
from tests import util

quiet = logging.getLogger('page-test')
quiet.setLevel(logging.CRITICAL)


def _oracle(stateful):
    wd, maparr, mapping = util.small_model()
    o = OracleRater(util.to_model(wd), mapping, 8, logger=quiet)
    o.stateful = stateful
    o.incremental = not stateful
    return o, maparr


def test_mets_context():
    assert pg.mets_get_context('http://kant_aufklaerung_1784') == [179]
    assert pg.mets_get_context('http://x/blumenbach_anatomie_1805') == [181]
    assert pg.mets_get_context('no-year') == [0]
    assert pg.mets_get_context(None) == [0]


@pytest.mark.parametrize('level,n_edges', [('region', 3), ('line', 11), ('word', 59), ('glyph', None)])
def test_linear_graph_shapes(level, n_edges):
    _, maparr = _oracle(True)
    doc = pg.PageDoc.fromstring(synth.page_xml(3, glyphs=True, mapping_arr=maparr))
    graph, start, end = pg.page_get_linear_graph_at(level, doc, logger=quiet)
    edges = [graph.edges[u, v] for u, v in zip(range(start, end), range(start + 1, end + 1))]
    assert start == 0 and end == graph.number_of_edges()
    if n_edges is not None:
        assert end == n_edges                       # 2 regions x 3 lines x 5 words, separators in between
    spaces = [e['alternatives'][0].Unicode for e in edges if e['element'] is None]
    if level == 'word':
        assert spaces.count('\n') == 5 and spaces.count(' ') == 24
        for e in edges:
            if e['element'] is not None:            # 5 alternatives in the file, 4 survive the filter
                assert len(e['alternatives']) == 4
                assert all(isinstance(a.conf, float) for a in e['alternatives'])
                assert e['alternatives'][0].conf - e['alternatives'][-1].conf < pg.CHOICE_THRESHOLD_CONF
    if level == 'glyph':
        text = ''.join(e['alternatives'][0].Unicode for e in edges)
        words = [pg.page_element_unicode0(w) for r in doc.get_TextRegion() for l in r.get_TextLine() for w in l.get_Word()]
        assert text.replace('\n', ' ').split(' ') == words
    if level in ('region', 'line'):
        assert all(s == '\n' for s in spaces)


def _check_consistent(doc, level):
    for region in doc.get_TextRegion():
        lines = region.get_TextLine()
        for line in lines:
            words = line.get_Word()
            for word in words:
                assert len(word.get_TextEquiv()) == 1
                if level == 'glyph':
                    glyphs = word.get_Glyph()
                    assert all(len(g.get_TextEquiv()) == 1 for g in glyphs)
                    assert pg.page_element_unicode0(word) == ''.join(pg.page_element_unicode0(g) for g in glyphs)
                    assert math.isclose(pg.page_element_conf0(word),
                                        sum(pg.page_element_conf0(g) for g in glyphs) / len(glyphs), rel_tol=1e-9)
            assert pg.page_element_unicode0(line) == ' '.join(pg.page_element_unicode0(w) for w in words)
            assert math.isclose(pg.page_element_conf0(line), sum(pg.page_element_conf0(w) for w in words) / len(words),
                                rel_tol=1e-9)
        assert pg.page_element_unicode0(region) == '\n'.join(pg.page_element_unicode0(l) for l in lines)


def test_plain_rating_scores_first_choices():
    o, maparr = _oracle(True)
    xml = synth.page_xml(5, mapping_arr=maparr)
    doc = pg.PageDoc.fromstring(xml)
    first = [[pg.page_element_unicode0(w) for w in l.get_Word()] for r in doc.get_TextRegion() for l in r.get_TextLine()]
    text = '\n'.join(' '.join(ws) for ws in first)
    pr = pg.PageRater(o, textequiv_level='word', alternative_decoding=False, lm_weight=0.3, identifier='a_b_1784',
                      logger=quiet)
    assert pr.process(doc) == [doc]
    o2, _ = _oracle(True)
    probs = o2.rate(text, [179])
    words = [w for r in doc.get_TextRegion() for l in r.get_TextLine() for w in l.get_Word()]
    i = 0
    for w, t in zip(words, [t for ws in first for t in ws]):
        tes = w.get_TextEquiv()
        assert len(tes) == 1 and tes[0].Unicode == t
        want = 0.3 * (sum(probs[i:i + len(t)]) / len(t)) + 0.7 * 0.90
        assert math.isclose(float(tes[0].conf), want, rel_tol=1e-6)
        i += len(t) + 1
    assert len(pr.stats) == 1 and pr.stats[0][2] > 1.0
    # the document survives a round trip through the serialiser
    again = pg.PageDoc.fromstring(doc.tostring())
    assert [pg.page_element_unicode0(w) for r in again.get_TextRegion() for l in r.get_TextLine() for w in l.get_Word()] == \
        [t for ws in first for t in ws]


@pytest.mark.parametrize('level', ['word', 'glyph'])
def test_alternative_decoding_chains_pages_and_updates_levels(level):
    o, maparr = _oracle(False)
    docs = [pg.PageDoc.fromstring(synth.page_xml(10 + k, page_id='p%d' % k, glyphs=(level == 'glyph'), n_regions=1,
                                                 lines_per_region=2, words_per_line=3, mapping_arr=maparr))
            for k in range(3)]
    allowed = [[[te.Unicode for te in w.get_TextEquiv()[:4]] for r in d.get_TextRegion() for l in r.get_TextLine()
                for w in l.get_Word()] for d in docs]
    pr = pg.PageRater(o, textequiv_level=level, alternative_decoding=True, beam_width=5, lm_weight=0.5, logger=quiet)
    finished = []
    for d in docs:
        finished += pr.process(d)
    assert finished == docs[:2]                     # a page is decided while the next one is searched
    finished += pr.finish()
    assert finished == docs and pr.finish() == []
    assert [s[0] for s in pr.stats] == ['p0', 'p1', 'p2']
    for d, alts in zip(docs, allowed):
        _check_consistent(d, level)
        if level == 'word':
            words = [w for r in d.get_TextRegion() for l in r.get_TextLine() for w in l.get_Word()]
            for w, a in zip(words, alts):
                te = w.get_TextEquiv()[0]
                assert te.Unicode in a and 0.0 < float(te.conf) <= 1.0


@pytest.mark.gpu
@pytest.mark.parametrize('alternative_decoding', [False, True])
def test_cuda_path_decides_like_the_oracle(alternative_decoding):
    from ocrd_keraslm_b200.lib import Rater
    wd, maparr, mapping = util.small_model()
    o = OracleRater(util.to_model(wd), mapping, 8, logger=quiet)
    o.stateful, o.incremental = (not alternative_decoding), alternative_decoding
    r = Rater.from_weights(wd, maparr, 8, stateful=not alternative_decoding, incremental=alternative_decoding, logger=quiet)
    results = []
    for rater in (o, r):
        docs = [pg.PageDoc.fromstring(synth.page_xml(20 + k, page_id='p%d' % k, n_regions=1, lines_per_region=2,
                                                     words_per_line=4, mapping_arr=maparr)) for k in range(2)]
        pg.rate_pages(rater, docs, textequiv_level='word', alternative_decoding=alternative_decoding, beam_width=7,
                      lm_weight=0.5, identifier='x_y_1790', logger=quiet)
        results.append([(w.id, w.get_TextEquiv()[0].Unicode, float(w.get_TextEquiv()[0].conf))
                        for d in docs for reg in d.get_TextRegion() for l in reg.get_TextLine() for w in l.get_Word()])
    assert [(a, b) for a, b, _ in results[0]] == [(a, b) for a, b, _ in results[1]]
    np.testing.assert_allclose([c for _, _, c in results[0]], [c for _, _, c in results[1]], rtol=2e-3, atol=1e-5)
