"""PAGE-XML -> lattice -> PAGE-XML glue around the rating path, without the ocrd package.

What the OCR-D processor of the reference does on either side of ``Rater.rate`` / ``Rater.rate_best``
(/root/reference/ocrd_keraslm/wrapper/rate.py): read the text annotation of a PAGE file at one hierarchy level,
build the linear graph of TextEquiv alternatives with the implicit white space between tokens (:343-424,
:612-644), rate it, write the chosen alternatives and their scores back (:426-441, :321-330) and make the levels
above consistent (:478-590).  The real deployment keeps using the reference's ``KerasRate`` processor unchanged on
top of this package's ``Rater``; this module exists so that the path can be run and measured on
PAGE files where ocrd is not installed (benchmarks, tests), on ``xml.etree`` only.

Not reproduced: the tokenisation repair that needs ocrd's PageValidator (:592-610, :646-660) -- white space is
always inserted by the plain rules --, and ReadingOrder groups (document order is used).
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

    def _children(self, tag):
        return [Segment(self.doc, n) for n in self.node.findall(self.doc.q(tag))]

    def get_TextRegion(self):
        return self._children('TextRegion')

    def get_TextLine(self):
        return self._children('TextLine')

    def get_Word(self):
        return self._children('Word')

    def get_Glyph(self):
        return self._children('Glyph')

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


def _filter_choices(textequivs):
    """At most CHOICE_THRESHOLD_NUM alternatives, none more than CHOICE_THRESHOLD_CONF below the first (:657-672)."""
    if not textequivs:
        return []
    textequivs = textequivs[:min(CHOICE_THRESHOLD_NUM, len(textequivs))]
    for te in textequivs:
        te.set_conf(float(te.conf) if te.conf else 1.0)
    conf0 = textequivs[0].conf
    return [te for te in textequivs if conf0 - te.conf < CHOICE_THRESHOLD_CONF]


def _add_element(graph, start_node, element, textequivs):
    graph.add_node(start_node + 1)
    graph.add_edge(start_node, start_node + 1, element=element, alternatives=_filter_choices(textequivs))
    return start_node + 1


def _add_space(graph, start_node, space):
    """Pseudo-element edge for the implicit white space between two tokens (:623-644, without the repair)."""
    return _add_element(graph, start_node, None, [TextEquiv(Unicode=space, conf=1.0)])


def _get_edges(graph, start_node):
    return [graph.edges[in_, out] for in_, out in nx.bfs_edges(graph, start_node)]


def page_get_linear_graph_at(level, doc, logger=None):
    """Linear graph of the page's elements at `level`: one edge per element with its (filtered) alternatives, plus
    one edge per implicit white space: a newline between lines and regions, a space between words (:343-424).
    Returns (graph, start node, end node)."""
    assert level in LEVELS
    logger = logger or logging.getLogger(__name__)
    graph = nx.DiGraph(level=level)
    graph.add_node(0)
    start_node = 0
    regions = doc.get_TextRegion()
    if not regions:
        logger.warning("Page contains no text regions")
    page_start_node = start_node
    first_region = True
    for region in regions:
        if level == 'region':
            textequivs = region.get_TextEquiv()
            if not first_region:
                start_node = _add_space(graph, start_node, '\n')
            if textequivs:
                start_node = _add_element(graph, start_node, region, textequivs)
            else:
                logger.warning("Region '%s' contains no text results", region.id)
            first_region = False
            continue
        lines = region.get_TextLine()
        if not lines:
            logger.warning("Region '%s' contains no text lines", region.id)
        first_line = True
        for line in lines:
            if level == 'line':
                textequivs = line.get_TextEquiv()
                if not first_line or not first_region:
                    start_node = _add_space(graph, start_node, '\n')
                if textequivs:
                    start_node = _add_element(graph, start_node, line, textequivs)
                else:
                    logger.warning("Line '%s' contains no text results", line.id)
                first_line = False
                continue
            words = line.get_Word()
            if not words:
                logger.warning("Line '%s' contains no words", line.id)
            first_word = True
            for word in words:
                textequivs = word.get_TextEquiv()
                if not first_word or not first_line or not first_region:
                    start_node = _add_space(graph, start_node, '\n' if first_word else ' ')
                if level == 'word':
                    if textequivs:
                        start_node = _add_element(graph, start_node, word, textequivs)
                    else:
                        logger.warning("Word '%s' contains no text results", word.id)
                    first_word = False
                    continue
                glyphs = word.get_Glyph()
                if not glyphs:
                    logger.warning("Word '%s' contains no glyphs", word.id)
                for glyph in glyphs:
                    textequivs = glyph.get_TextEquiv()
                    if textequivs:
                        start_node = _add_element(graph, start_node, glyph, textequivs)
                    else:
                        logger.warning("Glyph '%s' contains no text results", glyph.id)
                first_word = False
            first_line = False
        first_region = False
    return graph, page_start_node, start_node


def page_update_from_path(level, path, entropy, logger=None):
    """Apply a decided path: keep only the chosen alternative on every element, with its score as confidence
    (:426-441).  Returns (average probability, character perplexity, element perplexity)."""
    logger = logger or logging.getLogger(__name__)
    strlen = 0
    for element, textequiv, score in path:
        if element:                                 # not just space
            textequiv.set_conf(score)
            element.set_TextEquiv([textequiv])      # delete others
            strlen += len(textequiv.Unicode)
        else:
            strlen += 1
    ent = entropy / strlen
    avg = pow(2.0, -ent)
    ppl = pow(2.0, ent)                             # character level
    ppll = pow(2.0, ent * strlen / len(path))       # textequiv level (including spaces/newlines)
    logger.info("avg: %.3f, char ppl: %.3f, %s ppl: %.3f", avg, ppl, level, ppll)
    return avg, ppl, ppll


def page_element_unicode0(element):
    tes = element.get_TextEquiv()
    return (tes[0].Unicode or '') if tes else ''


def page_element_conf0(element):
    tes = element.get_TextEquiv()
    return float(tes[0].conf or "1.0") if tes else 1.0


def page_update_higher_textequiv_levels(level, doc, overwrite=True):
    """Make the TextEquivs of all hierarchy levels above `level` consistent with it: concatenate the first
    alternatives (glyphs directly, words by space, lines and regions by newline, nothing where a ``join`` Relation
    links two neighbours), average their confidences; right-to-left / bottom-to-top orders are honoured (:478-590)."""
    if level == 'region':
        return
    joins = doc.joins()
    page = doc.page
    for region in doc.get_AllTextRegions():
        subregions = region.get_TextRegion()
        if subregions:                              # already visited: concatenate them
            region_unicode = page_element_unicode0(subregions[0])
            for sub, nxt in zip(subregions, subregions[1:]):
                if (sub.id, nxt.id) not in joins:
                    region_unicode += '\n'
                region_unicode += page_element_unicode0(nxt)
            region_conf = sum(page_element_conf0(sub) for sub in subregions) / len(subregions)
        else:
            lines = region.get_TextLine()
            if (region.get('textLineOrder') or page.get('textLineOrder')) == 'bottom-to-top':
                lines = list(reversed(lines))
            if level != 'line':
                for line in lines:
                    words = line.get_Word()
                    if (line.get('readingDirection') or region.get('readingDirection') or
                            page.get('readingDirection')) == 'right-to-left':
                        words = list(reversed(words))
                    if level != 'word':
                        for word in words:
                            glyphs = word.get_Glyph()
                            if (word.get('readingDirection') or line.get('readingDirection') or
                                    region.get('readingDirection') or page.get('readingDirection')) == 'right-to-left':
                                glyphs = list(reversed(glyphs))
                            word_unicode = ''.join(page_element_unicode0(glyph) for glyph in glyphs)
                            word_conf = sum(page_element_conf0(glyph) for glyph in glyphs)
                            if glyphs:
                                word_conf /= len(glyphs)
                            if not word.get_TextEquiv() or overwrite:
                                word.set_TextEquiv([TextEquiv(Unicode=word_unicode, conf=word_conf)])
                    line_unicode = ' '.join(page_element_unicode0(word) for word in words)
                    line_conf = sum(page_element_conf0(word) for word in words)
                    if words:
                        line_conf /= len(words)
                    if not line.get_TextEquiv() or overwrite:
                        line.set_TextEquiv([TextEquiv(Unicode=line_unicode, conf=line_conf)])
            region_unicode = ''
            region_conf = 0
            if lines:
                region_unicode = page_element_unicode0(lines[0])
                for line, next_line in zip(lines, lines[1:]):
                    words, next_words = line.get_Word(), next_line.get_Word()
                    if not (words and next_words and (words[-1].id, next_words[0].id) in joins):
                        region_unicode += '\n'
                    region_unicode += page_element_unicode0(next_line)
                region_conf = sum(page_element_conf0(line) for line in lines) / len(lines)
        if not region.get_TextEquiv() or overwrite:
            region.set_TextEquiv([TextEquiv(Unicode=region_unicode, conf=region_conf)])


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
