"""Region-adjacency graph of a superpixel label map (oracle; test infrastructure only).

PARITY UNPINNED: the reference calls ``skimage.graph.rag_mean_color(image_for_slic, segments)``
(models/region_graph/extract_rg_embeddings.py:216) and reads only ``rag.edges()`` (:220).
scikit-image (>=0.21.0, unpinned, models/region_graph/requirements.txt:4) is not installed, so this
restates its published algorithm: ``graph.RAG(labels, connectivity=2)`` runs
``scipy.ndimage.generic_filter`` over the label image with the 3x3 footprint of
``generate_binary_structure(2, 2)``, ``mode='nearest'``, and its callback adds an undirected
networkx edge ``(centre, v)`` for every footprint value ``v != centre`` not yet present.
scikit-image zeroes the "backward" half of that footprint before filtering (each unordered pixel pair
is then visited once, from its raster-earlier pixel); this restatement keeps the FULL 3x3 footprint.
The edge SET is identical -- a pair of touching labels is found from either side -- only networkx's
insertion order can differ, and everything downstream compares on the canonical sorted order.
"""
from __future__ import annotations

import numpy as np


def rag_edges_networkx(segments: np.ndarray):
    """Edges in the reference's own order (networkx insertion order), as a list of (a, b) label pairs.

    Slow (one Python callback per pixel): use on small label maps or a few images.
    """
    import networkx as nx
    from scipy import ndimage as ndi

    g = nx.Graph()

    def visit(values, graph):
        vals = values.astype(int)
        c = vals[len(vals) // 2]
        for v in vals:
            if v != c and not graph.has_edge(c, v):
                graph.add_edge(c, v)
        return 0.0

    fp = ndi.generate_binary_structure(segments.ndim, 2)
    sink = np.broadcast_to(1.0, segments.shape).copy()
    ndi.generic_filter(segments, function=visit, footprint=fp, mode="nearest", output=sink,
                       extra_arguments=(g,))
    return [(int(a), int(b)) for a, b in g.edges()]


def rag_edge_set(segments: np.ndarray) -> np.ndarray:
    """Same edge set by shifted-array compares: sorted unique ``[M,2]`` int64 pairs with a < b.

    Two labels are adjacent iff they occupy two pixels that touch in the 8-neighbourhood; comparing
    every pixel with its E, S, SE and SW neighbours visits every unordered pixel pair exactly once.
    ``mode='nearest'`` border replication only ever compares a pixel with itself or an in-image
    neighbour, so it adds nothing.
    """
    s = np.asarray(segments).astype(np.int64)
    pairs = []
    for a, b in ((s[:, :-1], s[:, 1:]), (s[:-1, :], s[1:, :]), (s[:-1, :-1], s[1:, 1:]), (s[:-1, 1:], s[1:, :-1])):
        m = a != b
        if m.any():
            pairs.append(np.stack([np.minimum(a[m], b[m]), np.maximum(a[m], b[m])], 1))
    if not pairs:
        return np.zeros((0, 2), dtype=np.int64)
    return np.unique(np.concatenate(pairs, 0), axis=0)


def compact_ids(segments: np.ndarray):
    """Label -> node id map (extract_rg_embeddings.py:146-161,210-211).

    The reference walks ``region_id in range(segments.max() + 1)`` and numbers the labels whose mask is
    non-empty 0..N-1 in ascending label order; labels outside ``[0, max]`` (negative) are never visited.
    Returns ``(id_of_label [max+1] int64 with -1 for absent labels, present labels ascending)``.
    """
    s = np.asarray(segments)
    top = int(s.max())
    id_of = np.full(max(top + 1, 0), -1, dtype=np.int64)
    present = np.unique(s[s >= 0]).astype(np.int64)
    id_of[present] = np.arange(len(present))
    return id_of, present


def directed_edge_index(segments: np.ndarray) -> np.ndarray:
    """Canonical ``edge_index`` ``[2,E]`` int64: both directions of every RAG edge whose two labels are
    mapped (extract_rg_embeddings.py:220-225), columns sorted lexicographically by (row 0, row 1).

    The reference's own column order is networkx insertion order; parity is defined on this sorted form
    (SURVEY.md §7 "Edge order").
    """
    id_of, _ = compact_ids(segments)
    und = rag_edge_set(segments)
    if len(und):
        ok = (und[:, 0] >= 0) & (und[:, 1] >= 0)
        und = und[ok]
    if len(und) == 0:
        return np.zeros((2, 0), dtype=np.int64)
    a, b = id_of[und[:, 0]], id_of[und[:, 1]]
    ei = np.concatenate([np.stack([a, b]), np.stack([b, a])], 1)
    order = np.lexsort((ei[1], ei[0]))
    return ei[:, order]


def canonical_sort(edge_index: np.ndarray, *edge_values: np.ndarray):
    """Sort the columns of an ``edge_index`` (and matching per-edge arrays) lexicographically."""
    ei = np.asarray(edge_index)
    order = np.lexsort((ei[1], ei[0]))
    return (ei[:, order],) + tuple(np.asarray(v)[order] for v in edge_values)
