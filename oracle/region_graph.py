"""Region-graph construction from a label map (oracle; test infrastructure only).

Restates ``create_region_graph`` (models/region_graph/extract_rg_embeddings.py:138-246; identical
copies in test.py:90-182 and train.py:89-221) from the label map onward.  SLIC and Canny are upstream
of the hot path: the label map and the boolean edge map are inputs here.
"""
from __future__ import annotations

import numpy as np
from scipy import ndimage

from . import rag as _rag

GRAY_WEIGHTS = (0.2989, 0.5870, 0.1140)  # extract_rg_embeddings.py:151
NORM_SIDE = 256.0                       # hard-coded normaliser, :172-174


def _region_row(image, gray, canny, m, win=None):
    """The 15 features of one region whose boolean mask is ``m`` (:159-206), all in float64.

    ``win`` is an optional (y0, y1, x0, x1) crop that contains the region's bounding box grown by two
    pixels; morphology on the crop equals morphology on the full frame because scipy's
    ``binary_dilation`` uses border value 0 and a 2-step cross dilation never reaches further than two
    pixels.  Positions are always measured in full-frame coordinates.
    """
    y0, x0 = (0, 0) if win is None else (win[0], win[2])
    px = image[m]
    mean_rgb = px.mean(axis=0)
    std_rgb = px.std(axis=0)                        # population std (ddof=0), :165
    g = gray[m]
    coords = np.argwhere(m)
    cy = (coords[:, 0].mean() + y0) / NORM_SIDE     # :172
    cx = (coords[:, 1].mean() + x0) / NORM_SIDE     # :173
    size = len(px) / (NORM_SIDE * NORM_SIDE)        # :174
    perimeter = np.sum(ndimage.binary_dilation(m) ^ m)          # 4-connected cross, :177
    area = m.sum()
    compact = (perimeter ** 2) / (4 * np.pi * area + 1e-10)     # :179
    edge_density = canny[m].mean()                               # :182
    ring = ndimage.binary_dilation(m, iterations=2) & ~m         # L1 distance <= 2, :185-186
    contrast = 0.0
    if ring.any():
        contrast = np.linalg.norm(mean_rgb - image[ring].mean(axis=0))  # :190
    row = np.concatenate([mean_rgb, std_rgb, [g.mean()], [g.std()], [cx, cy], [size], [compact],
                          [contrast], [edge_density], [np.var(g)]])      # layout :195-206
    return np.nan_to_num(row, nan=0.0)              # :207


def node_features(image, segments, canny, faithful=False):
    """``x`` ``[N,15]`` float32 and the label -> node id map.

    ``faithful=True`` walks full-frame boolean masks exactly as the reference loop does (:154-211,
    seconds per image); the default crops every region to its bounding box grown by two pixels first,
    which is arithmetically identical (same pixels, same order) and ~100x faster.
    """
    image = np.asarray(image, dtype=np.float64)
    seg = np.asarray(segments)
    canny = np.asarray(canny).astype(bool)
    gray = np.dot(image[..., :3], GRAY_WEIGHTS)     # :151
    hh, ww = seg.shape
    id_of, present = _rag.compact_ids(seg)
    rows = []
    boxes = ndimage.find_objects(np.where(seg >= 0, seg + 1, 0).astype(np.int64)) if not faithful else None
    for lab in present:
        if faithful:
            m = seg == lab
            rows.append(_region_row(image, gray, canny, m))
        else:
            sl = boxes[lab]
            ya, yb = max(sl[0].start - 2, 0), min(sl[0].stop + 2, hh)
            xa, xb = max(sl[1].start - 2, 0), min(sl[1].stop + 2, ww)
            m = seg[ya:yb, xa:xb] == lab
            rows.append(_region_row(image[ya:yb, xa:xb], gray[ya:yb, xa:xb], canny[ya:yb, xa:xb], m,
                                    win=(ya, yb, xa, xb)))
    x = np.array(rows).astype(np.float32) if rows else np.zeros((0, 15), np.float32)   # FloatTensor, :213
    return x, id_of


def edge_weights(x: np.ndarray, edge_index: np.ndarray) -> np.ndarray:
    """Gaussian-kernel similarity per directed edge (:227-235), all-float32 arithmetic.

    The third factor reads feature 12, which is the boundary contrast although the reference names the
    variable ``edge_diff`` (layout :195-206); the oracle follows the code.
    """
    if edge_index.shape[1] == 0:
        return np.zeros((0,), np.float32)
    xi, xj = x[edge_index[0]], x[edge_index[1]]
    d = (xi[:, :3] - xj[:, :3]).astype(np.float32)
    color = np.sqrt((d * d).sum(axis=1, dtype=np.float32), dtype=np.float32)
    tex = np.abs(xi[:, 6] - xj[:, 6]).astype(np.float32)
    con = np.abs(xi[:, 12] - xj[:, 12]).astype(np.float32)
    f32 = np.float32
    w = np.exp(-color / f32(0.15), dtype=f32) * np.exp(-tex / f32(0.08), dtype=f32) * np.exp(-con / f32(0.1), dtype=f32)
    return w.astype(np.float32)


def build_region_graph(image, segments, canny, faithful=False):
    """(x [N,15] f32, edge_index [2,E] i64 canonically sorted, edge_attr [E,1] f32, id_of)."""
    x, id_of = node_features(image, segments, canny, faithful=faithful)
    ei = _rag.directed_edge_index(np.asarray(segments))
    w = edge_weights(x, ei)
    return x, ei, w.reshape(-1, 1), id_of


def region_labels(segments, mask, instance_mask, edge_mask):
    """Per-region training labels (train.py:169-176): mean of the target over the region, thresholds
    0.5 / 0.5 / 0.3.  Targets are float arrays in [0,1] (``ToTensor`` of an 8-bit PNG)."""
    seg = np.asarray(segments)
    _, present = _rag.compact_ids(seg)
    y, iy, ey = [], [], []
    for lab in present:
        m = seg == lab
        y.append(1 if mask[m].mean() > 0.5 else 0)
        iy.append(1 if instance_mask[m].mean() > 0.5 else 0)
        ey.append(1.0 if edge_mask[m].mean() > 0.3 else 0.0)
    return np.array(y, np.int64), np.array(iy, np.int64), np.array(ey, np.float32)
