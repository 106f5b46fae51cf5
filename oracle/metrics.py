"""Pixel-mask paint and segmentation metrics (oracle; test infrastructure only).

Restates ``utils/metrics.py:9-92`` (PINNED: checked against the reference file executed verbatim, see
``oracle/make_golden.py``) and the paint loop of ``models/region_graph/test.py:238-244``.
"""
from __future__ import annotations

import numpy as np

EPS = 1e-8
KEYS = ("iou", "dice", "precision", "recall", "f1", "mae", "accuracy")


def paint_mask(mask_logits: np.ndarray, segments: np.ndarray, label_offset_fix: bool = False,
               id_of_label: np.ndarray | None = None) -> np.ndarray:
    """``[H,W]`` float64 probability mask (test.py:238-244).

    p = softmax(mask_out, 1)[:, 1] in float32; every pixel whose label is ``l`` receives ``p[l]`` when
    ``0 <= l < len(p)``, else 0.  The reference indexes ``p`` by the RAW label, not by the compacted node
    id, so with labels 1..L label ``l`` shows node ``l``'s probability and label ``L`` stays 0 (SURVEY.md
    §8 A8); ``label_offset_fix=True`` paints ``p[id_of_label[l]]`` instead.
    """
    lg = np.asarray(mask_logits, dtype=np.float32)
    z = lg - lg.max(axis=1, keepdims=True)
    e = np.exp(z, dtype=np.float32)
    p = (e / e.sum(axis=1, keepdims=True, dtype=np.float32))[:, 1].astype(np.float32) if len(lg) else np.zeros(0, np.float32)
    seg = np.asarray(segments)
    out = np.zeros(seg.shape, dtype=np.float64)
    if label_offset_fix:
        idx = np.where((seg >= 0) & (seg < len(id_of_label)), id_of_label[np.clip(seg, 0, len(id_of_label) - 1)], -1)
    else:
        idx = seg
    ok = (idx >= 0) & (idx < len(p))
    out[ok] = p[idx[ok]]
    return out


def confusion_counts(pred, gt, threshold=0.5):
    """Integer (tp, sum_pred, sum_gt, n_equal, n) with strict ``>`` thresholds (metrics.py:11-12)."""
    pb = np.asarray(pred) > threshold
    gb = np.asarray(gt) > threshold
    return int((pb & gb).sum()), int(pb.sum()), int(gb.sum()), int((pb == gb).sum()), int(pb.size)


def metrics_from_counts(tp, sp, sg, neq, n, abs_err_sum):
    """The seven values of ``evaluate_segmentation`` (metrics.py:57-76) from exact integer counts and the
    sum of |pred - gt|.

    The reference forms the five ratios in FLOAT32: the 0/1 masks are ``astype(np.float32)`` (:11-12), their
    sums are float32 scalars (exact integers below 2^24), and ``float32_scalar + 1e-8`` stays float32, so the
    ``1e-8`` guard only matters when the denominator is otherwise 0.  MAE and accuracy are float64.
    """
    f = np.float32
    tp, sp, sg = f(tp), f(sp), f(sg)
    eps = f(1e-8)
    fp, fn = sp - tp, sg - tp                              # :36-37 (sums of exact 0/1 products)
    iou = tp / ((sp + sg - tp) + eps)                      # :14-17
    dice = (f(2) * tp) / (sp + sg + eps)                   # :26
    prec = tp / (tp + fp + eps)                            # :39
    rec = tp / (tp + fn + eps)                             # :40
    f1 = f(2) * (prec * rec) / (prec + rec + eps)          # :41
    return {"iou": float(iou), "dice": float(dice), "precision": float(prec), "recall": float(rec), "f1": float(f1),
            "mae": abs_err_sum / n, "accuracy": neq / n}


def evaluate_segmentation(pred, gt, threshold=0.5):
    pred = np.asarray(pred)
    gt = np.asarray(gt)
    tp, sp, sg, neq, n = confusion_counts(pred, gt, threshold)
    return metrics_from_counts(tp, sp, sg, neq, n, float(np.abs(pred - gt).sum()))


F32_KEYS = ("iou", "dice", "precision", "recall", "f1")


def batch_evaluate(preds, gts, threshold=0.5):
    """Mean and population std per key over images (metrics.py:78-92).

    The five ratio metrics are np.float32 scalars in the reference, so ``np.mean`` / ``np.std`` of the list
    run with a float32 accumulator (numpy pairwise order); MAE and accuracy are float64.  The device path
    accumulates float64 partial sums instead (order-free across ranks): it agrees to ~1e-7, the float32
    rounding of the reference's own accumulator -- tests compare batch statistics at 1e-6.
    """
    per = [evaluate_segmentation(p, g, threshold) for p, g in zip(preds, gts)]
    out = {}
    for k in KEYS:
        v = np.array([m[k] for m in per], dtype=np.float32 if k in F32_KEYS else np.float64)
        out[k] = float(v.mean())
        out[f"{k}_std"] = float(v.std())
    return out


def image_level_f1(pred_cls, true_cls):
    """``calculate_f1_score`` (models/multimodal/train_multimodal.py:197-220): binary confusion counts
    over image-level class ids, per-class precision/recall/F1 with +1e-8 in each ratio, macro average."""
    pred_cls = np.asarray(pred_cls)
    true_cls = np.asarray(true_cls)
    tp = int(((pred_cls == 1) & (true_cls == 1)).sum())
    fp = int(((pred_cls == 1) & (true_cls == 0)).sum())
    fn = int(((pred_cls == 0) & (true_cls == 1)).sum())
    tn = int(((pred_cls == 0) & (true_cls == 0)).sum())
    return f1_from_confusion(tp, fp, fn, tn)


def f1_from_confusion(tp, fp, fn, tn):
    p1 = tp / (tp + fp + EPS)
    r1 = tp / (tp + fn + EPS)
    f1_1 = 2 * (p1 * r1) / (p1 + r1 + EPS)
    p0 = tn / (tn + fn + EPS)
    r0 = tn / (tn + fp + EPS)
    f1_0 = 2 * (p0 * r0) / (p0 + r0 + EPS)
    return {"f1_class_0": f1_0, "f1_class_1": f1_1, "f1_avg": (f1_0 + f1_1) / 2,
            "precision_1": p1, "recall_1": r1, "counts": (tp, fp, fn, tn)}
