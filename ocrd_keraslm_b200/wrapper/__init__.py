"""Callers of the rating path: the ocrd-free PAGE-XML glue."""
from .page import PageDoc, PageRater, rate_pages, page_get_linear_graph_at, page_update_higher_textequiv_levels

__all__ = ['PageDoc', 'PageRater', 'rate_pages', 'page_get_linear_graph_at', 'page_update_higher_textequiv_levels']
