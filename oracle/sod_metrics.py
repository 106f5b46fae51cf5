"""S-measure, E-measure, F-measure (beta^2 = 0.3) and weighted F-measure (oracle; test infrastructure only).

PARITY UNPINNED -- NOT IN THE REFERENCE.  ``utils/metrics.py:9-92`` stops at IoU / Dice / P / R / F1 / MAE / accuracy
(SURVEY F3); BASELINE.json's north_star nevertheless names "MAE, S/E/F-measure" with a ``|dS| <= 1e-3`` tolerance, so
this module restates the PUBLISHED definitions, in float64 numpy, in the form the camouflaged-object-detection
literature evaluates them (the per-threshold curves on the 8-bit quantised prediction, the adaptive threshold
``min(2 * mean(pred), 1)``, dataset-level max / mean taken over the dataset-averaged curve):

  S-measure   Fan et al., "Structure-measure: a new way to evaluate foreground maps", ICCV 2017 (alpha = 0.5):
              object-aware term (mean / sample-std of the prediction over GT foreground and of 1 - prediction over GT
              background) and region-aware term (SSIM of the four quadrants around the GT centroid, area-weighted).
  E-measure   Fan et al., "Enhanced-alignment measure for binary foreground map evaluation", IJCAI 2018: for a binary
              prediction the enhanced alignment matrix takes four values (one per confusion cell), so E is a closed form
              of (TP, FP, FN, TN); curve over 256 thresholds + adaptive threshold.
  F-measure   (1 + b2) P R / (b2 P + R) with b2 = 0.3 (Achanta et al., CVPR 2009): curve + adaptive threshold.
  weighted F  Margolin et al., "How to evaluate foreground maps?", CVPR 2014 (beta^2 = 1): pixel dependency (error of a
              background pixel := error of its nearest foreground pixel, 7x7 Gaussian sigma 5, min with the raw error
              inside the GT) and pixel importance (2 - exp(ln(0.5) / 5 * distance to the GT)).  The nearest-foreground
              tie rule is not part of the definition; here: the lexicographically smallest (y, x) among the nearest
              (tests compare with scipy's distance_transform_edt indices as a sanity bound).

Conventions shared with the device kernels (csrc/sod_metrics.cu): ``pred`` is a float map in [0, 1] used as it is (no
min-max stretch: it is a painted probability, not an 8-bit saliency image); GT foreground = ``gt > 0.5`` like
utils/metrics.py:12 (byte >= 128 for a u8 mask / 255); 8-bit quantisation ``q = floor(pred * 255)`` (``astype(uint8)``).
"""
from __future__ import annotations

import numpy as np

EPS = np.spacing(1)            # Matlab eps, as the published evaluation code uses
KEYS = ("s_measure", "e_adaptive", "e_mean", "e_max", "f_adaptive", "f_mean", "f_max", "wf_measure")
BETA2 = 0.3


# ------------------------------------------------------------------------------------------------ S-measure
def _s_object(x_vals):
    """2 x / (x^2 + 1 + sigma + eps) over the selected pixels (sample std, ddof = 1)."""
    n = x_vals.size
    x = x_vals.mean()
    sigma = np.sqrt(((x_vals - x) ** 2).sum() / (n - 1)) if n > 1 else 0.0     # a 1-pixel class has no sample std: 0
    return 2 * x / (x * x + 1 + sigma + EPS)


def _ssim(p, g):
    n = p.size
    nm1 = max(n - 1, 1)                 # a 1-pixel quadrant: the published code divides 0 by 0 here; use 0 variance
    x, y = p.mean(), g.mean()
    sx = ((p - x) ** 2).sum() / nm1
    sy = ((g - y) ** 2).sum() / nm1
    sxy = ((p - x) * (g - y)).sum() / nm1
    alpha = 4 * x * y * sxy
    beta = (x * x + y * y) * (sx + sy)
    if alpha != 0:
        return alpha / (beta + EPS)
    return 1.0 if beta == 0 else 0.0


def s_measure(pred, gt, alpha=0.5):
    pred = np.asarray(pred, np.float64)
    g = np.asarray(gt) > 0.5
    h, w = g.shape
    y = g.mean()
    if y == 0:
        return float(1 - pred.mean())
    if y == 1:
        return float(pred.mean())
    gf = g.astype(np.float64)
    so = y * _s_object(pred[g]) + (1 - y) * _s_object(1 - pred[~g])
    cy, cx = np.argwhere(g).mean(axis=0).round()           # numpy round: half to even
    cx, cy = int(cx) + 1, int(cy) + 1
    area = h * w
    w1, w2, w3 = cx * cy / area, cy * (w - cx) / area, (h - cy) * cx / area
    w4 = 1 - w1 - w2 - w3
    quads = ((slice(0, cy), slice(0, cx)), (slice(0, cy), slice(cx, w)), (slice(cy, h), slice(0, cx)), (slice(cy, h), slice(cx, w)))
    sr = 0.0
    for wt, (ys, xs) in zip((w1, w2, w3, w4), quads):
        if pred[ys, xs].size == 0:      # a centroid on the last row / column leaves an empty quadrant of weight 0
            continue
        sr += wt * _ssim(pred[ys, xs], gf[ys, xs])
    return float(max(0.0, alpha * so + (1 - alpha) * sr))


# ------------------------------------------------------------------------------------------------ E / F from counts
def e_from_counts(tp, fp, n_fg, n):
    """Enhanced-alignment measure of a binary prediction with ``tp`` foreground hits and ``fp`` false alarms
    (arrays allowed) against a GT with ``n_fg`` of ``n`` pixels set."""
    tp = np.asarray(tp, np.float64)
    fp = np.asarray(fp, np.float64)
    pf = tp + fp                      # predicted foreground
    pb = n - pf
    if n_fg == 0:
        s = pb
    elif n_fg == n:
        s = pf
    else:
        fn = n_fg - tp
        tn = pb - fn
        mp, mg = pf / n, n_fg / n
        s = np.zeros_like(pf)
        for cnt, a, b in ((tp, 1 - mp, 1 - mg), (fp, 1 - mp, 0 - mg), (fn, 0 - mp, 1 - mg), (tn, 0 - mp, 0 - mg)):
            align = 2 * (a * b) / (a * a + b * b + EPS)
            s = s + (align + 1) ** 2 / 4 * cnt
    return s / (n - 1 + EPS)


def f_from_counts(tp, fp, n_fg, beta2=BETA2):
    tp = np.asarray(tp, np.float64)
    ps = np.asarray(tp + fp, np.float64).copy()
    ps[ps == 0] = 1
    prec = tp / ps
    rec = tp / max(n_fg, 1)
    den = beta2 * prec + rec
    den = np.where(den == 0, 1.0, den)
    return (1 + beta2) * prec * rec / den


def curves(pred, gt):
    """(F curve [256], E curve [256]); entry i is the threshold ``q >= 255 - i`` on the 8-bit quantised prediction."""
    pred = np.asarray(pred, np.float64)
    g = np.asarray(gt) > 0.5
    q = (pred * 255).astype(np.uint8)
    fg_hist = np.bincount(q[g], minlength=256)
    bg_hist = np.bincount(q[~g], minlength=256)
    tp = np.cumsum(fg_hist[::-1])
    fp = np.cumsum(bg_hist[::-1])
    n_fg = int(g.sum())
    return f_from_counts(tp, fp, n_fg), e_from_counts(tp, fp, n_fg, g.size)


def adaptive(pred, gt):
    """(F, E) at the adaptive threshold ``min(2 * mean(pred), 1)`` (``pred >= threshold``)."""
    pred = np.asarray(pred, np.float64)
    g = np.asarray(gt) > 0.5
    thr = min(2 * pred.mean(), 1.0)
    b = pred >= thr
    tp, fp, n_fg = int((b & g).sum()), int((b & ~g).sum()), int(g.sum())
    f = 0.0 if tp == 0 else float(f_from_counts(np.array([tp]), np.array([fp]), n_fg)[0])
    return f, float(e_from_counts(np.array([tp]), np.array([fp]), n_fg, g.size)[0])


# ------------------------------------------------------------------------------------------------ weighted F
def gauss7():
    """Matlab ``fspecial('gaussian', 7, 5)``."""
    ax = np.arange(-3, 4, dtype=np.float64)
    k = np.exp(-(ax[:, None] ** 2 + ax[None, :] ** 2) / (2 * 5.0 * 5.0))
    k[k < np.finfo(np.float64).eps * k.max()] = 0
    return k / k.sum()


def nearest_foreground(g):
    """Exact Euclidean feature transform of a boolean map: (squared distance [H,W] int64, nearest y, nearest x), ties to
    the lexicographically smallest (y, x).  Column pass (nearest set row per column, upper one on ties) then row pass."""
    h, w = g.shape
    big = 1 << 40
    ys = np.arange(h)
    col_d = np.full((h, w), big, np.int64)
    col_y = np.zeros((h, w), np.int64)
    for x in range(w):
        rows = ys[g[:, x]]
        if rows.size == 0:
            continue
        d = np.abs(ys[:, None] - rows[None, :])
        j = d.argmin(axis=1)                       # first minimum = smallest row
        col_d[:, x] = d[ys, j]
        col_y[:, x] = rows[j]
    xs = np.arange(w)
    d2 = np.full((h, w), big, np.int64)
    ny = np.zeros((h, w), np.int64)
    nx = np.zeros((h, w), np.int64)
    for y in range(h):
        cd = col_d[y]
        tot = np.where(cd[None, :] >= big, big, (xs[:, None] - xs[None, :]) ** 2 + cd[None, :] ** 2)    # [x, x']
        best = tot.min(axis=1)
        cand = tot == best[:, None]
        key = np.where(cand, col_y[y][None, :] * w + xs[None, :], np.iinfo(np.int64).max)
        j = key.argmin(axis=1)
        d2[y], ny[y], nx[y] = best, col_y[y][j], xs[j]
    return d2, ny, nx


def weighted_f(pred, gt, beta2=1.0):
    pred = np.asarray(pred, np.float64)
    g = np.asarray(gt) > 0.5
    if not g.any():
        return 0.0
    from scipy.ndimage import convolve
    gf = g.astype(np.float64)
    d2, ny, nx = nearest_foreground(g)
    dst = np.sqrt(d2.astype(np.float64))
    e = np.abs(pred - gf)
    et = e.copy()
    et[~g] = e[ny[~g], nx[~g]]
    ea = convolve(et, weights=gauss7(), mode="constant", cval=0)
    min_e_ea = np.where(g & (ea < e), ea, e)
    b = np.where(~g, 2 - np.exp(np.log(0.5) / 5 * dst), 1.0)
    ew = min_e_ea * b
    tpw = gf.sum() - ew[g].sum()
    fpw = ew[~g].sum()
    r = 1 - ew[g].mean()
    p = tpw / (tpw + fpw + EPS)
    return float((1 + beta2) * r * p / (r + beta2 * p + EPS))


# ------------------------------------------------------------------------------------------------ bundles
def evaluate_sod(pred, gt):
    """The eight per-image values (KEYS) plus the two curves."""
    fc, ec = curves(pred, gt)
    fa, ea = adaptive(pred, gt)
    vals = (s_measure(pred, gt), ea, float(ec.mean()), float(ec.max()), fa, float(fc.mean()), float(fc.max()), weighted_f(pred, gt))
    return dict(zip(KEYS, vals)), fc, ec


def dataset_summary(preds, gts):
    """Dataset-level numbers the way the literature reports them: mean of the per-image S / adaptive / weighted-F values,
    and max / mean of the dataset-AVERAGED F and E curves."""
    per, fcs, ecs = zip(*(evaluate_sod(p, g) for p, g in zip(preds, gts)))
    fc, ec = np.mean(fcs, axis=0), np.mean(ecs, axis=0)
    out = {k: float(np.mean([m[k] for m in per])) for k in ("s_measure", "e_adaptive", "f_adaptive", "wf_measure")}
    out.update({"f_max": float(fc.max()), "f_mean": float(fc.mean()), "e_max": float(ec.max()), "e_mean": float(ec.mean())})
    return out
