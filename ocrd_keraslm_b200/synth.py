"""Synthetic vocabulary, text, lattices and random-init models.

There is no network for datasets or trained checkpoints, so benchmarks and tests
use seeded synthetic inputs of the shapes BASELINE.json names.  Weight
initialisers restate the Keras defaults used by the reference's
``Rater.configure`` (/root/reference/ocrd_keraslm/lib/rating.py:104-114, 139-145):
glorot-uniform kernels, orthogonal recurrent kernels, zero bias with unit forget
bias, N(0, 0.001) embeddings; a "trained-like" variant scales things up so that
numerical differences become visible in parity tests.
"""
import numpy as np

UNMAPPED_CHAR = u'☃'


def vocabulary():
    """255 mapped code points; index 0 = unmapped.  Returns uint32[256] as stored in /config/mapping."""
    cps = [0x0A] + list(range(0x20, 0x7F)) + list(range(0xA1, 0x140))
    assert len(cps) == 255
    return np.array([0] + cps, dtype=np.uint32)


def mapping_from_array(arr):
    """(char->index, index->char) exactly as rating.py:961-963 derives it."""
    c_i = dict((chr(int(c)), i) for i, c in enumerate(arr) if c > 0)
    i_c = dict((i, chr(int(c))) for i, c in enumerate(arr) if c > 0)
    return c_i, i_c


def text(seed, size, mapping_arr=None, unmapped_rate=0.01):
    """Zipf-like synthetic text with word and line breaks and a few unmapped chars."""
    mapping_arr = vocabulary() if mapping_arr is None else mapping_arr
    rng = np.random.default_rng(seed)
    cps = np.array([c for c in mapping_arr[1:] if c not in (0x0A, 0x20)], dtype=np.uint32)
    p = 1.0 / (np.arange(len(cps)) + 5.0)
    p /= p.sum()
    out = rng.choice(cps, size=size, p=p)
    # word breaks every 2..9 chars, line breaks every 40..80 chars
    pos = 0
    gaps = rng.integers(3, 11, size=size // 3 + 2)
    idx = np.cumsum(gaps)
    idx = idx[idx < size]
    out[idx] = 0x20
    lgaps = rng.integers(40, 81, size=size // 40 + 2)
    lidx = np.cumsum(lgaps)
    lidx = lidx[lidx < size]
    out[lidx] = 0x0A
    if unmapped_rate > 0:
        mask = rng.random(size) < unmapped_rate
        out[mask] = ord(UNMAPPED_CHAR)
    del pos
    return out.astype('<u4').tobytes().decode('utf-32-le')


def _orthogonal(rng, rows, cols):
    a = rng.standard_normal((max(rows, cols), min(rows, cols)))
    q, r = np.linalg.qr(a)
    q *= np.sign(np.diag(r))
    if rows < cols:
        q = q.T
    return q[:rows, :cols]


def weights(seed, width, depth, voc_size=256, contexts=((200, 10),), style='keras', dtype=np.float32):
    """Random-init weights as a dict in the reference's naming (Appendix A of SURVEY.md).

    style='keras'   : Keras default initialisers (near-uniform predictions)
    style='trained' : scaled so gates saturate and predictions are peaked
    """
    rng = np.random.default_rng(seed)
    W = width
    out = {}
    if style == 'keras':
        out['char_embedding'] = rng.normal(0, 0.001, (voc_size, W))
        for k, (rows, dim) in enumerate(contexts):
            out['context%d_embedding' % (k + 1)] = rng.normal(0, 0.001, (rows, dim))
    else:
        out['char_embedding'] = rng.normal(0, 1.0 / np.sqrt(W), (voc_size, W)) * 3.0
        for k, (rows, dim) in enumerate(contexts):
            out['context%d_embedding' % (k + 1)] = rng.normal(0, 0.5, (rows, dim))
    din = W + sum(d for _, d in contexts)
    for l in range(depth):
        d_l = din if l == 0 else W
        limit = np.sqrt(6.0 / (d_l + 4 * W))
        K = rng.uniform(-limit, limit, (d_l, 4 * W))
        R = np.concatenate([_orthogonal(rng, W, W) for _ in range(4)], axis=1) if W <= 256 \
            else _orthogonal(rng, W, 4 * W)
        b = np.zeros(4 * W)
        b[W:2 * W] = 1.0
        if style != 'keras':
            K *= 1.5
            R *= 1.5
            b += rng.normal(0, 0.1, 4 * W)
        out['lstm_%d' % (l + 1)] = (K, R, b)
    res = {}
    for k, v in out.items():
        res[k] = tuple(np.ascontiguousarray(a, dtype=dtype) for a in v) if isinstance(v, tuple) \
            else np.ascontiguousarray(v, dtype=dtype)
    return res


class Alt(object):
    """Stand-in for ocrd_models TextEquivType: .Unicode, .conf, .index."""

    def __init__(self, text, conf, index=None):
        self.Unicode = text
        self.conf = conf
        self.index = index

    def __repr__(self):
        return 'Alt(%r, %.2f)' % (self.Unicode, self.conf)


class Element(object):
    """Stand-in for a PAGE element (needs only .id)."""

    def __init__(self, id_):
        self.id = id_


def lattice(seed, n_words=200, n_alts=5, mapping_arr=None, words_per_line=10, min_len=2, max_len=9):
    """Linear lattice of word alternatives like wrapper/rate.py:343-424 builds from a PAGE.

    Nodes 0..N; word edges carry `n_alts` alternatives (base + single-substitution variants,
    conf 0.90, 0.88, ...), separator edges carry a single ' ' (or '\\n' every
    `words_per_line` words) alternative with conf 1.0.  Returns (graph, start, end).
    """
    import networkx as nx
    mapping_arr = vocabulary() if mapping_arr is None else mapping_arr
    rng = np.random.default_rng(seed)
    letters = [chr(int(c)) for c in mapping_arr[1:] if c not in (0x0A, 0x20)]
    p = 1.0 / (np.arange(len(letters)) + 5.0)
    p /= p.sum()
    g = nx.DiGraph()
    node = 0
    g.add_node(node)
    for wi in range(n_words):
        n = int(rng.integers(min_len, max_len + 1))
        base = ''.join(rng.choice(letters, size=n, p=p))
        alts = [Alt(base, 0.90, 0)]
        for a in range(1, n_alts):
            pos = int(rng.integers(0, n))
            sub = letters[int(rng.integers(0, len(letters)))]
            alts.append(Alt(base[:pos] + sub + base[pos + 1:], 0.90 - 0.02 * a, a))
        g.add_edge(node, node + 1, element=Element('w%04d' % wi), alternatives=alts)
        node += 1
        sep = '\n' if (wi + 1) % words_per_line == 0 else ' '
        g.add_edge(node, node + 1, element=None, alternatives=[Alt(sep, 1.0, 0)])
        node += 1
    return g, 0, node


def page_xml(seed, page_id='p0001', n_regions=2, lines_per_region=3, words_per_line=5, n_alts=5, glyphs=False,
             mapping_arr=None, min_len=2, max_len=7):
    """Synthetic PAGE-XML (2019 namespace) with OCR alternatives, the input shape of wrapper/rate.py:343-424.

    Words carry `n_alts` TextEquiv alternatives (@index, @conf = 0.90, 0.88, ... : the first four survive the
    reference's choice filter, the drop stays below 0.1); with `glyphs`, every character of the first alternative
    is also a Glyph with two alternatives.  Lines and regions carry a (stale) single TextEquiv.
    """
    from xml.sax.saxutils import escape
    mapping_arr = vocabulary() if mapping_arr is None else mapping_arr
    rng = np.random.default_rng(seed)
    letters = [chr(int(c)) for c in mapping_arr[1:] if c not in (0x0A, 0x20)]
    p = 1.0 / (np.arange(len(letters)) + 5.0)
    p /= p.sum()
    ns = 'http://schema.primaresearch.org/PAGE/gts/pagecontent/2019-07-15'
    out = ['<?xml version="1.0" encoding="UTF-8"?>',
           '<PcGts xmlns="%s" pcGtsId="%s">' % (ns, page_id),
           '<Metadata><Creator>synth</Creator><Created>2026-01-01T00:00:00</Created>'
           '<LastChange>2026-01-01T00:00:00</LastChange></Metadata>',
           '<Page imageFilename="%s.png" imageWidth="1000" imageHeight="1000">' % page_id]

    def te(text, conf, index=None):
        return '<TextEquiv%s conf="%.2f"><Unicode>%s</Unicode></TextEquiv>' % (
            ' index="%d"' % index if index is not None else '', conf, escape(text))
    coords = '<Coords points="0,0 10,0 10,10 0,10"/>'
    for r in range(n_regions):
        out.append('<TextRegion id="r%d">%s' % (r, coords))
        for l in range(lines_per_region):
            out.append('<TextLine id="r%d_l%d">%s' % (r, l, coords))
            for w in range(words_per_line):
                n = int(rng.integers(min_len, max_len + 1))
                base = ''.join(rng.choice(letters, size=n, p=p))
                wid = 'r%d_l%d_w%d' % (r, l, w)
                out.append('<Word id="%s">%s' % (wid, coords))
                if glyphs:
                    for g, ch in enumerate(base):
                        other = letters[int(rng.integers(0, len(letters)))]
                        out.append('<Glyph id="%s_g%d">%s%s%s</Glyph>' % (wid, g, coords, te(ch, 0.95, 0),
                                                                           te(other, 0.90, 1)))
                for a in range(n_alts):
                    if a == 0:
                        alt = base
                    else:
                        pos = int(rng.integers(0, n))
                        alt = base[:pos] + letters[int(rng.integers(0, len(letters)))] + base[pos + 1:]
                    out.append(te(alt, 0.90 - 0.02 * a, a))
                out.append('</Word>')
            out.append(te('stale line text', 0.5))
            out.append('</TextLine>')
        out.append(te('stale region text', 0.5))
        out.append('</TextRegion>')
    out.append('</Page></PcGts>')
    return '\n'.join(out)
