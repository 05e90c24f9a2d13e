"""PAGE-XML -> lattice -> PAGE-XML glue around the rating path, without the ocrd package.

What the OCR-D processor of the reference does on either side of ``Rater.rate`` / ``Rater.rate_best``
(/root/reference/ocrd_keraslm/wrapper/rate.py): read the text annotation of a PAGE file at one hierarchy level,
build the linear graph of TextEquiv alternatives with the implicit white space between tokens (:343-424,
:612-644), rate it, write the chosen alternatives and their scores back (:426-441, :321-330) and make the levels
above consistent (:478-590).  The real deployment keeps using the reference's ``KerasRate`` processor unchanged on
top of this package's ``Rater``; this module exists so that the path can be run and measured on
PAGE files where ocrd is not installed (benchmarks, tests), on ``xml.etree`` only.

Not reproduced: the tokenisation repair that needs ocrd's PageValidator (:592-610, :646-660) -- white space is
always inserted by the plain rules.  The hierarchy is walked generically over a table of levels (one recursive walk
builds the graph, one bottom-up pass rebuilds the higher levels), ``ReadingOrder`` is honoured where the reference
honours it (the order of nested regions, :510-531).
"""
import logging
import math
import re
import xml.etree.ElementTree as ET

import networkx as nx

CHOICE_THRESHOLD_NUM = 4      # maximum number of choices to try per element (reference :41)
CHOICE_THRESHOLD_CONF = 0.1   # maximum score drop from best choice to try per element (:42)
BEAM_CLUSTERING_ENABLE = True  # (:44)
BEAM_CLUSTERING_DIST = 5      # (:45)

LEVELS = ('region', 'line', 'word', 'glyph')


class TextEquiv(object):
    """One text alternative of a segment: what the reference reads of ocrd's TextEquivType."""

    __slots__ = ('Unicode', 'conf', 'index', 'node')

    def __init__(self, Unicode='', conf=None, index=None, node=None):
        self.Unicode = Unicode
        self.conf = conf
        self.index = index
        self.node = node            # the xml element this was read from (None: made here)

    def set_conf(self, conf):
        self.conf = conf

    def __repr__(self):
        return 'TextEquiv(%r, conf=%r, index=%r)' % (self.Unicode, self.conf, self.index)


class Segment(object):
    """A TextRegion / TextLine / Word / Glyph element with the accessors the rating code uses."""

    CHILD = {'TextRegion': 'TextLine', 'TextLine': 'Word', 'Word': 'Glyph', 'Glyph': None}

    def __init__(self, doc, node):
        self.doc = doc
        self.node = node
        self.id = node.get('id')

    def children(self, tag):
        """Direct child segments of one kind ('TextRegion', 'TextLine', 'Word', 'Glyph'), in document order."""
        return [Segment(self.doc, n) for n in self.node.findall(self.doc.q(tag))]

    def get_TextRegion(self):
        return self.children('TextRegion')

    def get_TextLine(self):
        return self.children('TextLine')

    def get_Word(self):
        return self.children('Word')

    def get_Glyph(self):
        return self.children('Glyph')

    def get_TextEquiv(self):
        """Alternatives in document order (OCR engines write them best first, with @index)."""
        out = []
        for n in self.node.findall(self.doc.q('TextEquiv')):
            u = n.find(self.doc.q('Unicode'))
            idx = n.get('index')
            out.append(TextEquiv(u.text or '' if u is not None else '', n.get('conf'),
                                 int(idx) if idx is not None else None, n))
        return out

    def set_TextEquiv(self, textequivs):
        """Replace all TextEquiv children by the given ones (kept behind the other children, as the schema wants;
        TextStyle, if any, stays last)."""
        for n in self.node.findall(self.doc.q('TextEquiv')):
            self.node.remove(n)
        style = self.node.find(self.doc.q('TextStyle'))
        pos = list(self.node).index(style) if style is not None else len(self.node)
        for te in textequivs:
            n = ET.Element(self.doc.q('TextEquiv'))
            if te.index is not None:
                n.set('index', str(te.index))
            if te.conf is not None:
                n.set('conf', repr(float(te.conf)))
            u = ET.SubElement(n, self.doc.q('Unicode'))
            u.text = te.Unicode
            self.node.insert(pos, n)
            pos += 1
            te.node = n

    def get(self, attr):
        return self.node.get(attr)


class PageDoc(object):
    """A parsed PAGE-XML file (any namespace version)."""

    def __init__(self, tree):
        self.tree = tree
        self.root = tree.getroot()
        m = re.match(r'\{(.*)\}', self.root.tag)
        self.ns = m.group(1) if m else ''
        if self.ns:
            ET.register_namespace('', self.ns)
        self.page = self.root.find(self.q('Page'))
        if self.page is None:
            raise ValueError('not a PAGE-XML document: no Page element')

    @classmethod
    def parse(cls, source):
        """From a file name or file object."""
        return cls(ET.parse(source))

    @classmethod
    def fromstring(cls, text):
        return cls(ET.ElementTree(ET.fromstring(text)))

    def q(self, tag):
        return '{%s}%s' % (self.ns, tag) if self.ns else tag

    def get_pcGtsId(self):
        return self.root.get('pcGtsId')

    def get_TextRegion(self):
        return [Segment(self, n) for n in self.page.findall(self.q('TextRegion'))]

    def get_AllTextRegions(self):
        """All text regions, nested ones before the regions that contain them (depth first, :520-527)."""
        out = []

        def walk(node):
            for n in node.findall(self.q('TextRegion')):
                walk(n)
                out.append(Segment(self, n))
        walk(self.page)
        return out

    def joins(self):
        """(source id, target id) of every Relation of type join (:500-509)."""
        out = []
        rel = self.page.find(self.q('Relations'))
        if rel is not None:
            for r in rel.findall(self.q('Relation')):
                if r.get('type') == 'join':
                    s, t = r.find(self.q('SourceRegionRef')), r.find(self.q('TargetRegionRef'))
                    if s is not None and t is not None:
                        out.append((s.get('regionRef'), t.get('regionRef')))
        return out

    def reading_order(self):
        """{region id: (kind of the ReadingOrder entry, its @index or None)} for every reference below Page/ReadingOrder,
        groups included and followed recursively (reference :457-476)."""
        out = {}

        def collect(group):
            for entry in group:
                kind = entry.tag.rsplit('}', 1)[-1]
                if kind not in ('RegionRef', 'RegionRefIndexed', 'OrderedGroup', 'OrderedGroupIndexed', 'UnorderedGroup',
                                'UnorderedGroupIndexed'):
                    continue
                index = entry.get('index')
                out[entry.get('regionRef')] = (kind, int(index) if index is not None else None)
                if 'Group' in kind:
                    collect(entry)

        ro = self.page.find(self.q('ReadingOrder'))
        if ro is not None:
            for top in ro:
                collect(top)
        return out

    def tostring(self):
        return ET.tostring(self.root, encoding='unicode')

    def write(self, target):
        self.tree.write(target, encoding='utf-8', xml_declaration=True)


def mets_get_context(identifier):
    """Context variables from the document identifier: ``.../author_title_year`` -> [ceil(year / 10)], else [0]
    (reference :332-341)."""
    context = [0]
    if identifier:
        name = identifier.split('/')[-1]
        year = name.split('_')[-1]
        if year.isnumeric():
            context = [math.ceil(int(year) / 10)]
    return context


# ---------------------------------------------------------------------------------------------------------------
# The PAGE hierarchy as data: everything below walks these tables instead of spelling the four levels out.
_SEGMENT_TAGS = ('TextRegion', 'TextLine', 'Word', 'Glyph')            # depth 0..3
_LEVEL_DEPTH = dict((name, depth) for depth, name in enumerate(LEVELS))
_NOUNS = ('Region', 'Line', 'Word', 'Glyph')                           # for the log messages
_PARTS = ('text regions', 'text lines', 'words', 'glyphs')
_WORD_DEPTH = 2
# what joins the first texts of a segment's children into its own text, and the attribute that can reverse their order
_JOINER = ('\n', ' ', '')                                               # region <- lines, line <- words, word <- glyphs
_ORDER_ATTR = (('textLineOrder', 'bottom-to-top'), ('readingDirection', 'right-to-left'),
               ('readingDirection', 'right-to-left'))


def _filter_choices(textequivs):
    """The alternatives worth searching (reference :657-672): the first CHOICE_THRESHOLD_NUM of the element, confidences
    as floats (missing = 1.0), minus those that fall CHOICE_THRESHOLD_CONF or more below the first."""
    kept = []
    for te in textequivs[:CHOICE_THRESHOLD_NUM]:
        te.set_conf(float(te.conf) if te.conf else 1.0)
        if kept and kept[0].conf - te.conf >= CHOICE_THRESHOLD_CONF:
            continue
        kept.append(te)
    return kept


def _reading_sequence(doc, target, warn):
    """The page as the language model reads it, for elements at depth `target`: yields (separator, None) for the implicit
    white space and (None, segment) for the elements, in document order.

    White space hangs on regions (region level), lines (line level) or words (word and glyph level): every such carrier
    except the one whose sibling indices are all zero is preceded by a newline -- or by a blank, if it is a word that is
    not the first of its line (reference :343-424: the ``first_*`` flags there are positional, so a page whose first
    region is empty starts with a separator; kept)."""
    carrier = min(target, _WORD_DEPTH)

    def descend(segment, depth, path):
        if depth == carrier and any(path):
            yield ' ' if depth == _WORD_DEPTH and path[-1] else '\n', None
        if depth == target:
            yield None, segment
            return
        children = segment.children(_SEGMENT_TAGS[depth + 1])
        if not children:
            warn("%s '%s' contains no %s", _NOUNS[depth], segment.id, _PARTS[depth + 1])
        for i, child in enumerate(children):
            for item in descend(child, depth + 1, path + (i,)):
                yield item

    regions = doc.get_TextRegion()
    if not regions:
        warn("Page contains no %s", _PARTS[0])
    for i, region in enumerate(regions):
        for item in descend(region, 0, (i,)):
            yield item


def page_get_linear_graph_at(level, doc, logger=None):
    """Linear graph of the page at `level` (reference :343-424): nodes 0..n, one edge per element that has text results
    (attributes ``element`` and ``alternatives``: its filtered TextEquivs) and one per implicit white space (``element``
    None, a single alternative with confidence 1).  Returns (graph, start node, end node)."""
    assert level in LEVELS
    logger = logger or logging.getLogger(__name__)
    graph = nx.DiGraph(level=level)
    graph.add_node(0)
    end = 0
    for separator, segment in _reading_sequence(doc, _LEVEL_DEPTH[level], logger.warning):
        if segment is None:
            alternatives = [TextEquiv(Unicode=separator, conf=1.0)]
        else:
            alternatives = _filter_choices(segment.get_TextEquiv())
            if not alternatives:
                logger.warning("%s '%s' contains no text results", _NOUNS[_LEVEL_DEPTH[level]], segment.id)
                continue
        graph.add_edge(end, end + 1, element=segment, alternatives=alternatives)
        end += 1
    return graph, 0, end


def _get_edges(graph, start_node):
    """Edge attribute dicts of a linear graph, in path order."""
    out = []
    node = start_node
    while True:
        successors = list(graph.successors(node))
        if not successors:
            return out
        out.append(graph.edges[node, successors[0]])
        node = successors[0]


def page_update_from_path(level, path, entropy, logger=None):
    """Write a decided path [(element, alternative, score)] back (reference :426-441): every element keeps the chosen
    alternative only, with the score as its confidence.  Returns (average probability, character perplexity,
    perplexity per element incl. white space) from the path's entropy in bits."""
    logger = logger or logging.getLogger(__name__)
    n_chars = 0
    for element, textequiv, score in path:
        if element is None:                         # implicit white space: one character, nothing to write
            n_chars += 1
            continue
        textequiv.set_conf(score)
        element.set_TextEquiv([textequiv])
        n_chars += len(textequiv.Unicode)
    bits_per_char = entropy / n_chars
    stats = (2.0 ** -bits_per_char, 2.0 ** bits_per_char, 2.0 ** (entropy / len(path)))
    logger.info("avg: %.3f, char ppl: %.3f, %s ppl: %.3f", stats[0], stats[1], level, stats[2])
    return stats


def page_element_unicode0(element):
    """Text of the element's first alternative ('' if it has none)."""
    for te in element.get_TextEquiv():
        return te.Unicode or ''
    return ''


def page_element_conf0(element):
    """Confidence of the element's first alternative (1.0 if missing)."""
    for te in element.get_TextEquiv():
        return float(te.conf) if te.conf else 1.0
    return 1.0


def _first_texts(segments):
    return [page_element_unicode0(seg) for seg in segments]


def _mean_conf(segments):
    return sum(page_element_conf0(seg) for seg in segments) / len(segments) if segments else 0


def _glue(texts, glued):
    """Concatenate with the level's joiner except where `glued[i]` says that texts i and i+1 belong together."""
    out = texts[0]
    for text, (joiner, skip) in zip(texts[1:], glued):
        out += text if skip else joiner + text
    return out


def page_update_higher_textequiv_levels(level, doc, overwrite=True):
    """Make the text of every hierarchy level above `level` the concatenation of the level below it (reference :478-590).

    Bottom-up over the tables above: a word is its glyphs' first texts, a line its words joined by blanks, a region its
    lines joined by newlines; confidences are averaged.  Children are taken in the order their ``readingDirection`` /
    ``textLineOrder`` (own, else inherited from the enclosing segments and the page) asks for; regions that contain
    regions are the newline-joined texts of those, in ``ReadingOrder`` where the page defines one for all of them, and
    are visited after them; a ``Relation`` of type join between two neighbours (regions, or the last and first word of
    two lines) drops the newline.  `overwrite` False keeps text that is already there."""
    bottom = _LEVEL_DEPTH[level]
    if bottom == 0:
        return
    joins = set(doc.joins())
    order = doc.reading_order()

    def put(segment, text, conf):
        if overwrite or not segment.get_TextEquiv():
            segment.set_TextEquiv([TextEquiv(Unicode=text, conf=conf)])

    def ordered_children(segment, depth, inherited):
        children = segment.children(_SEGMENT_TAGS[depth + 1])
        attr, reverse_value = _ORDER_ATTR[depth]
        wanted = next((n.get(attr) for n in (segment,) + inherited if n.get(attr)), None)
        return list(reversed(children)) if wanted == reverse_value else children

    def rebuild(segment, depth, inherited):
        """Text and confidence of `segment` from its children, after rebuilding those (down to the rated level)."""
        children = ordered_children(segment, depth, inherited)
        if depth + 1 < bottom:
            for child in children:
                rebuild(child, depth + 1, (segment,) + inherited)
        if depth == 0:
            if not children:
                return put(segment, '', 0)
            # between two lines the newline goes if their adjoining words (in document order) are joined
            edge_words = [line.children('Word') for line in children]
            glued = [('\n', bool(a and b and (a[-1].id, b[0].id) in joins)) for a, b in zip(edge_words, edge_words[1:])]
        else:
            glued = [(_JOINER[depth], False)] * max(len(children) - 1, 0)
        put(segment, _glue(_first_texts(children), glued) if children else '', _mean_conf(children))

    for region in doc.get_AllTextRegions():             # nested regions come before the regions that contain them
        inner = region.get_TextRegion()
        if not inner:
            rebuild(region, 0, (doc.page,))
            continue
        ranks = [order.get(sub.id) for sub in inner]
        if all(r is not None and r[1] is not None for r in ranks) and ranks[0][0].startswith('OrderedGroup'):
            inner = [sub for _, sub in sorted(zip((r[1] for r in ranks), inner), key=lambda pair: pair[0])]
        glued = [('\n', (a.id, b.id) in joins) for a, b in zip(inner, inner[1:])]
        put(region, _glue(_first_texts(inner), glued), _mean_conf(inner))


class PageRater(object):
    """Rates the PAGE files of one document in reading order with a configured ``Rater``.

    `alternative_decoding` False: the first alternatives of every page are concatenated and rated with
    ``Rater.rate`` (state carried from page to page), each element gets ``lm_weight * mean LM probability +
    (1 - lm_weight) * OCR confidence`` (reference :287-318).
    `alternative_decoding` True: ``Rater.rate_best`` over the graph of alternatives; pages are chained through the
    traceback, the decision for page k is taken while page k+1 is searched and the last page is closed with
    ``next_path`` (:235-285, :222-241).  The rater must then be non-stateful and incremental (:88-90).
    """

    def __init__(self, rater, textequiv_level='glyph', alternative_decoding=True, beam_width=10, lm_weight=0.5,
                 identifier=None, logger=None):
        assert textequiv_level in LEVELS
        self.rater = rater
        self.level = textequiv_level
        self.alternative_decoding = alternative_decoding
        self.beam_width = beam_width
        self.lm_weight = lm_weight
        self.context = mets_get_context(identifier)
        self.logger = logger or logging.getLogger(__name__)
        self._prev = None                   # (traceback, doc) of the page still waiting for its decision
        self.stats = []                     # (pcGtsId, avg, char ppl, element ppl) per finished page

    def process(self, doc):
        """Feed the next page.  Returns the pages finished by this call (plain mode: this page; alternative decoding:
        the previous page, once its best path is decided)."""
        graph, start_node, end_node = page_get_linear_graph_at(self.level, doc, logger=self.logger)
        if not self.alternative_decoding:
            self._rate_first_choices(doc, graph)
            return [doc]
        self.logger.info("Rating %d elements including its alternatives", end_node - start_node)
        path, entropy, traceback = self.rater.rate_best(
            graph, start_node, end_node,
            start_traceback=self._prev and self._prev[0],
            context=self.context, lm_weight=self.lm_weight, beam_width=self.beam_width,
            beam_clustering_dist=BEAM_CLUSTERING_DIST if BEAM_CLUSTERING_ENABLE else 0)
        done = []
        if self._prev:
            done.append(self._close(self._prev[1], path, entropy))
        self._prev = (traceback, doc)
        return done

    def finish(self):
        """Close the last page (alternative decoding): lock into the current best path (:222-241)."""
        done = []
        if self._prev:
            traceback, doc = self._prev
            path, entropy, _ = self.rater.next_path(traceback[0], ([], traceback[1]))
            done.append(self._close(doc, path, entropy))
            self._prev = None
        return done

    def _close(self, doc, path, entropy):
        avg, ppl, ppll = page_update_from_path(self.level, path, entropy, logger=self.logger)
        page_update_higher_textequiv_levels(self.level, doc)
        self.stats.append((doc.get_pcGtsId(), avg, ppl, ppll))
        return doc

    def _rate_first_choices(self, doc, graph):
        text = [(edge['element'], edge['alternatives']) for edge in _get_edges(graph, 0)]
        text = [(element, textequivs) for element, textequivs in text if textequivs]
        textstring = ''.join(textequivs[0].Unicode for element, textequivs in text)
        self.logger.info("Rating %d elements with a total of %d characters", len(text), len(textstring))
        if not textstring:
            return
        confidences = self.rater.rate(textstring, self.context)
        i = 0
        for element, textequivs in text:
            textequiv = textequivs[0]               # 1st choice only
            n = len(textequiv.Unicode)
            conf = sum(confidences[i:i + n]) / n if n else 0.0      # mean probability
            textequiv.set_conf(conf * self.lm_weight + textequiv.conf * (1. - self.lm_weight))
            if element:
                element.set_TextEquiv([textequiv])  # delete others
            i += n
        if i != len(confidences):
            self.logger.critical("Input text length and output scores length are off by %d characters",
                                 i - len(confidences))
        avg = sum(confidences) / len(confidences)
        ent = sum(-math.log(max(p, 1e-99), 2) for p in confidences) / len(confidences)
        self.stats.append((doc.get_pcGtsId(), avg, pow(2.0, ent), pow(2.0, ent * len(confidences) / len(text))))
        self.logger.info("avg: %.3f, char ppl: %.3f, %s ppl: %.3f", avg, self.stats[-1][2], self.level,
                         self.stats[-1][3])


def rate_pages(rater, docs, **kwargs):
    """All pages of one document; returns the PageRater (its .stats hold the per-page perplexities)."""
    pr = PageRater(rater, **kwargs)
    for doc in docs:
        pr.process(doc)
    pr.finish()
    return pr
