"""Segmentation metrics on the B200 reduction kernels.

Same names, arguments and return keys as the reference's ``utils/metrics.py`` (:9-92): numpy arrays (or
tensors) in, Python floats / dicts out.  Confusion counts are exact integers accumulated on the device;
ratios are formed in float64 with the reference's ``+1e-8`` guards.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib

KEYS = ("iou", "dice", "precision", "recall", "f1", "mae", "accuracy")
# A9' (not in the reference; published definitions, see csrc/sod_metrics.cu / oracle/sod_metrics.py)
SOD_KEYS = ("s_measure", "e_adaptive", "e_mean", "e_max", "f_adaptive", "f_mean", "f_max", "wf_measure")
SOD_PARTIAL = 2 * len(SOD_KEYS) + 512


def _device():
    if not torch.cuda.is_available():
        raise _lib.CmError("no CUDA device: the B200 metrics path has no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def _as_batch(a, dev):
    """-> [B, P] float32 CUDA tensor (a single image becomes a batch of one)."""
    t = a if torch.is_tensor(a) else torch.as_tensor(np.asarray(a))
    t = t.to(device=dev, dtype=torch.float32)
    return t.reshape(1, -1)


def dense_counts(preds: torch.Tensor, gts: torch.Tensor, threshold: float = 0.5):
    """Device entry: preds/gts ``[B, P]`` f32 CUDA -> (counts ``[B,4]`` i64: tp, sum_pred, sum_gt, n_equal;
    abs_err ``[B]`` f64)."""
    lib = _lib.load()
    _lib.require_cuda(preds, "pred")
    _lib.require_cuda(gts, "gt")
    if preds.shape != gts.shape or preds.dim() != 2:
        raise ValueError(f"pred {tuple(preds.shape)} and gt {tuple(gts.shape)} must both be [B, P]")
    preds, gts = _lib.f32c(preds), _lib.f32c(gts)
    B, P = preds.shape
    counts = torch.empty((B, 4), dtype=torch.int64, device=preds.device)
    abs_err = torch.empty(B, dtype=torch.float64, device=preds.device)
    _lib.check(lib.cm_metrics_dense(_lib.ptr(preds), _lib.ptr(gts), B, P, float(threshold), _lib.ptr(counts),
                                    _lib.ptr(abs_err), _lib.stream_ptr(preds.device)), "cm_metrics_dense")
    return counts, abs_err


def finalize(counts: torch.Tensor, abs_err: torch.Tensor, pixels: int):
    """(per_image ``[B,7]`` f64, partial ``[15]`` f64 = n, sum m_k, sum m_k^2) on the device."""
    lib = _lib.load()
    B = counts.shape[0]
    per = torch.empty((B, 7), dtype=torch.float64, device=counts.device)
    part = torch.empty(15, dtype=torch.float64, device=counts.device)
    _lib.check(lib.cm_metrics_finalize(_lib.ptr(counts), _lib.ptr(abs_err), B, int(pixels), _lib.ptr(per), _lib.ptr(part),
                                       _lib.stream_ptr(counts.device)), "cm_metrics_finalize")
    return per, part


def summary_from_partial(partial) -> dict:
    """``batch_evaluate``'s 14 keys (mean and population std per metric, utils/metrics.py:86-90) from the
    15-element partial-sum vector (after any cross-rank all-reduce)."""
    v = np.asarray(partial.detach().cpu() if torch.is_tensor(partial) else partial, dtype=np.float64)
    n = v[0]
    out = {}
    for k, name in enumerate(KEYS):
        mean = v[1 + k] / n
        var = max(v[8 + k] / n - mean * mean, 0.0)
        out[name] = float(mean)
        out[f"{name}_std"] = float(np.sqrt(var))
    return out


def _eval_one(pred, gt, threshold):
    dev = _device()
    p, g = _as_batch(pred, dev), _as_batch(gt, dev)
    counts, abs_err = dense_counts(p, g, threshold)
    per, _ = finalize(counts, abs_err, p.shape[1])
    return per[0].tolist()


def calculate_iou(pred_mask, gt_mask, threshold=0.5):
    return _eval_one(pred_mask, gt_mask, threshold)[0]


def calculate_dice(pred_mask, gt_mask, threshold=0.5):
    return _eval_one(pred_mask, gt_mask, threshold)[1]


def calculate_precision_recall_f1(pred_mask, gt_mask, threshold=0.5):
    m = _eval_one(pred_mask, gt_mask, threshold)
    return m[2], m[3], m[4]


def calculate_mae(pred_mask, gt_mask):
    return _eval_one(pred_mask, gt_mask, 0.5)[5]


def calculate_accuracy(pred_mask, gt_mask, threshold=0.5):
    return _eval_one(pred_mask, gt_mask, threshold)[6]


def evaluate_segmentation(pred_mask, gt_mask, threshold=0.5):
    """utils/metrics.py:57-76 -> dict with the seven keys."""
    return dict(zip(KEYS, _eval_one(pred_mask, gt_mask, threshold)))


def batch_evaluate(pred_masks, gt_masks, threshold=0.5):
    """utils/metrics.py:78-92 -> per-key mean and ``<key>_std`` (population std) over the images."""
    dev = _device()
    if len(pred_masks) == 0:
        raise ValueError("batch_evaluate needs at least one image")
    p = torch.cat([_as_batch(x, dev) for x in pred_masks], 0)
    g = torch.cat([_as_batch(x, dev) for x in gt_masks], 0)
    counts, abs_err = dense_counts(p, g, threshold)
    _, part = finalize(counts, abs_err, p.shape[1])
    return summary_from_partial(part)


def class_confusion(mask_out: torch.Tensor, true_cls: torch.Tensor, counts4: torch.Tensor | None = None):
    """Image-level tp, fp, fn, tn (train_multimodal.py:197-220) accumulated into ``counts4`` (i64 [4])."""
    lib = _lib.load()
    _lib.require_cuda(mask_out, "mask_out")
    mask_out = _lib.f32c(mask_out)
    true_cls = true_cls.to(device=mask_out.device, dtype=torch.int64).contiguous()
    if counts4 is None:
        counts4 = torch.zeros(4, dtype=torch.int64, device=mask_out.device)
    _lib.check(lib.cm_class_confusion(_lib.ptr(mask_out), _lib.ptr(true_cls), mask_out.shape[0], mask_out.shape[1],
                                      _lib.ptr(counts4), _lib.stream_ptr(mask_out.device)), "cm_class_confusion")
    return counts4


def f1_from_confusion(tp, fp, fn, tn):
    """train_multimodal.py:206-220: per-class precision / recall / F1 (+1e-8 guards), macro average."""
    eps = 1e-8
    p1, r1 = tp / (tp + fp + eps), tp / (tp + fn + eps)
    f1_1 = 2 * (p1 * r1) / (p1 + r1 + eps)
    p0, r0 = tn / (tn + fn + eps), tn / (tn + fp + eps)
    f1_0 = 2 * (p0 * r0) / (p0 + r0 + eps)
    return {"f1_class_0": f1_0, "f1_class_1": f1_1, "f1_avg": (f1_0 + f1_1) / 2, "precision_1": p1, "recall_1": r1}


# ---------------------------------------------------------------------------------------------------- A9'
def sod_metrics(pred: torch.Tensor, gt_u8: torch.Tensor, per_image=None, partial=None, weighted_f: bool = True):
    """Device entry: ``pred`` ``[B,H,W]`` f32 in [0,1], ``gt_u8`` ``[B,H,W]`` uint8 (foreground = byte >= 128), CUDA ->
    (per_image ``[B,8]`` f64 in ``SOD_KEYS`` order, partial ``[528]`` f64: sums, sums of squares, summed F / E curves).
    ``weighted_f=False`` skips the weighted F-measure (NaN in its slot)."""
    from .region_graph import WORKSPACE
    lib = _lib.load()
    _lib.require_cuda(pred, "pred")
    _lib.require_cuda(gt_u8, "gt")
    if pred.dim() != 3 or pred.shape != gt_u8.shape or gt_u8.dtype != torch.uint8:
        raise ValueError(f"pred {tuple(pred.shape)} f32 and gt {tuple(gt_u8.shape)} uint8 must both be [B,H,W]")
    pred, gt_u8 = _lib.f32c(pred), gt_u8.contiguous()
    B, H, W = pred.shape
    dev = pred.device
    if per_image is None:
        per_image = torch.empty((B, len(SOD_KEYS)), dtype=torch.float64, device=dev)
    if partial is None:
        partial = torch.empty(SOD_PARTIAL, dtype=torch.float64, device=dev)
    ws_bytes = lib.cm_sod_workspace_bytes(B, H, W)
    ws = WORKSPACE.get(dev, "sod", ws_bytes)
    _lib.check(lib.cm_sod_metrics(_lib.ptr(pred), _lib.ptr(gt_u8), B, H, W, 1 if weighted_f else 0, _lib.ptr(ws), ws_bytes,
                                  _lib.ptr(per_image), _lib.ptr(partial), _lib.stream_ptr(dev)), "cm_sod_metrics")
    return per_image, partial


def sod_metrics_painted(labels, gt_u8, mask_prob, node_ptr, id_of_label=None, label_cap=0, per_image=None, partial=None,
                        weighted_f: bool = False):
    """The same reductions on the mask PAINTED from ``labels`` ``[B,H,W]`` (int32 / int16) and the per-node probabilities
    (``cm_paint_metrics``' gather, models/region_graph/test.py:241-244) without writing the mask."""
    from .region_graph import WORKSPACE
    lib = _lib.load()
    for t, n in ((labels, "labels"), (gt_u8, "gt"), (mask_prob, "mask_prob"), (node_ptr, "node_ptr")):
        _lib.require_cuda(t, n)
    if labels.dtype not in (torch.int32, torch.int16) or gt_u8.dtype != torch.uint8 or labels.shape != gt_u8.shape or labels.dim() != 3:
        raise ValueError("labels [B,H,W] int32 / int16 and gt [B,H,W] uint8 expected")
    B, H, W = labels.shape
    dev = labels.device
    if per_image is None:
        per_image = torch.empty((B, len(SOD_KEYS)), dtype=torch.float64, device=dev)
    if partial is None:
        partial = torch.empty(SOD_PARTIAL, dtype=torch.float64, device=dev)
    ws_bytes = lib.cm_sod_workspace_bytes(B, H, W)
    ws = WORKSPACE.get(dev, "sod", ws_bytes)
    p = _lib.ptr
    _lib.check(lib.cm_sod_metrics_painted(p(labels.contiguous()), labels.element_size(), p(gt_u8.contiguous()), B, H, W, p(_lib.f32c(mask_prob)),
                                          p(node_ptr), p(id_of_label), int(label_cap), 1 if weighted_f else 0, p(ws), ws_bytes, p(per_image),
                                          p(partial), _lib.stream_ptr(dev)), "cm_sod_metrics_painted")
    return per_image, partial


RT_MAX_LABEL_CAP = 6000


class RegionTable:
    """A8 + A9 + A9' of a painted mask through the region x ground-truth contingency table (``cm_region_table_*``): the
    pixel pass needs labels + gt only, so it runs next to the graph build; behind the model one small block per image turns
    table rows x node probabilities into the same counts / sums the pixel kernels (``cm_paint_metrics``,
    ``cm_sod_metrics_painted``) give.  Weighted F is not part of it (it needs the pixel-level distance transform).

    ``build(labels, gt)`` -- labels ``[B,H,W]`` int32 / int16, gt ``[B,H,W]`` uint8 or bit-packed ``[B,H*W/8]`` uint8;
    ``metrics(mask_prob, node_ptr, ...)`` -- after the model, on the same stream or one that waited for ``build``."""

    def __init__(self, label_cap: int, lane_key: str = "region_table"):
        if not 0 < int(label_cap) <= RT_MAX_LABEL_CAP:
            raise ValueError(f"label_cap must be in 1..{RT_MAX_LABEL_CAP}")
        self.label_cap = int(label_cap)
        self._key = lane_key
        self._built = None

    def build(self, labels, gt):
        from .region_graph import WORKSPACE
        lib = _lib.load()
        _lib.require_cuda(labels, "labels")
        _lib.require_cuda(gt, "gt")
        if labels.dtype not in (torch.int32, torch.int16) or labels.dim() != 3 or gt.dtype != torch.uint8:
            raise ValueError("labels [B,H,W] int32 / int16 and gt uint8 expected")
        B, H, W = labels.shape
        packed = gt.dim() == 2
        if packed and ((H * W) % 8 or tuple(gt.shape) != (B, H * W // 8)):
            raise ValueError(f"packed gt must be [B, H*W/8] uint8, got {tuple(gt.shape)}")
        if not packed and gt.shape != labels.shape:
            raise ValueError("gt [B,H,W] uint8 expected")
        labels, gt = labels.contiguous(), gt.contiguous()
        dev = labels.device
        ws_bytes = lib.cm_region_table_workspace_bytes(B, H, W, self.label_cap)
        ws = WORKSPACE.get(dev, self._key, ws_bytes)
        p = _lib.ptr
        _lib.check(lib.cm_region_table_build(p(labels), labels.element_size(), p(gt), 1 if packed else 0, B, H, W, self.label_cap, p(ws), ws_bytes,
                                             _lib.stream_ptr(dev)), "cm_region_table_build")
        self._built = (labels, None if packed else gt, ws, ws_bytes, (B, H, W))
        return self

    def metrics(self, mask_prob, node_ptr, id_of_label=None, threshold: float = 0.5, counts=None, abs_err=None, per_image=None, partial=None,
                per_image_sod=None, partial_sod=None):
        """-> dict(counts [B,4] i64, abs_err [B], per_image [B,7], partial [15], per_image_sod [B,8], partial_sod [528])."""
        if self._built is None:
            raise RuntimeError("RegionTable.metrics before build")
        lib = _lib.load()
        labels, gt_bytes, ws, ws_bytes, (B, H, W) = self._built
        dev = labels.device
        _lib.require_cuda(mask_prob, "mask_prob")
        _lib.require_cuda(node_ptr, "node_ptr")
        f64 = dict(dtype=torch.float64, device=dev)
        counts = torch.empty((B, 4), dtype=torch.int64, device=dev) if counts is None else counts
        abs_err = torch.empty(B, **f64) if abs_err is None else abs_err
        per_image = torch.empty((B, 7), **f64) if per_image is None else per_image
        partial = torch.empty(15, **f64) if partial is None else partial
        per_image_sod = torch.empty((B, len(SOD_KEYS)), **f64) if per_image_sod is None else per_image_sod
        partial_sod = torch.empty(SOD_PARTIAL, **f64) if partial_sod is None else partial_sod
        p = _lib.ptr
        _lib.check(lib.cm_region_table_metrics(p(labels), labels.element_size(), p(gt_bytes), B, H, W, p(_lib.f32c(mask_prob)), p(node_ptr),
                                               p(id_of_label), self.label_cap, float(threshold), p(ws), ws_bytes, p(counts), p(abs_err),
                                               p(per_image), p(partial), p(per_image_sod), p(partial_sod), _lib.stream_ptr(dev)),
                   "cm_region_table_metrics")
        return {"counts": counts, "abs_err": abs_err, "per_image": per_image, "partial": partial, "per_image_sod": per_image_sod,
                "partial_sod": partial_sod}


def _sod_inputs(pred_mask, gt_mask):
    dev = _device()
    p = pred_mask if torch.is_tensor(pred_mask) else torch.as_tensor(np.asarray(pred_mask))
    g = gt_mask if torch.is_tensor(gt_mask) else torch.as_tensor(np.asarray(gt_mask))
    if p.dim() == 2:
        p, g = p[None], g[None]
    p = p.to(device=dev, dtype=torch.float32)
    if g.dtype != torch.uint8:                       # float GT in [0,1] (np.array(PIL 'L') / 255.0): foreground = > 0.5
        g = (g.to(dev) > 0.5).to(torch.uint8) * 255
    return p, g.to(dev)


def evaluate_sod(pred_mask, gt_mask):
    """S- / E- / F- / weighted-F-measure of one ``[H,W]`` prediction in [0,1] -> dict with the eight ``SOD_KEYS``."""
    p, g = _sod_inputs(pred_mask, gt_mask)
    per, _ = sod_metrics(p, g)
    return dict(zip(SOD_KEYS, per[0].tolist()))


def calculate_s_measure(pred_mask, gt_mask):
    return evaluate_sod(pred_mask, gt_mask)["s_measure"]


def calculate_e_measure(pred_mask, gt_mask):
    """Adaptive-threshold E-measure of one image."""
    return evaluate_sod(pred_mask, gt_mask)["e_adaptive"]


def calculate_f_measure(pred_mask, gt_mask):
    """Adaptive-threshold F-measure (beta^2 = 0.3) of one image."""
    return evaluate_sod(pred_mask, gt_mask)["f_adaptive"]


def calculate_weighted_f_measure(pred_mask, gt_mask):
    return evaluate_sod(pred_mask, gt_mask)["wf_measure"]


def sod_summary_from_partial(partial, n_images) -> dict:
    """Dataset-level numbers from the 528-element partial vector (after any cross-rank all-reduce) the way the
    literature reports them: mean (and population std) of the per-image values, and ``f_max / f_mean / e_max / e_mean``
    taken over the dataset-AVERAGED threshold curves."""
    v = np.asarray(partial.detach().cpu() if torch.is_tensor(partial) else partial, dtype=np.float64)
    n = float(n_images)
    k = len(SOD_KEYS)
    out = {}
    for i, name in enumerate(SOD_KEYS):
        mean = v[i] / n
        out[name + "_per_image"] = float(mean)
        out[name + "_per_image_std"] = float(np.sqrt(max(v[k + i] / n - mean * mean, 0.0)))
    fc, ec = v[2 * k:2 * k + 256] / n, v[2 * k + 256:2 * k + 512] / n
    for name in ("s_measure", "e_adaptive", "f_adaptive", "wf_measure"):
        out[name] = out[name + "_per_image"]
    if np.isnan(out["wf_measure"]):          # the weighted F-measure was not asked for
        for name in [k for k in out if k.startswith("wf_measure")]:
            del out[name]
    out.update({"f_max": float(fc.max()), "f_mean": float(fc.mean()), "e_max": float(ec.max()), "e_mean": float(ec.mean())})
    return out


def batch_evaluate_sod(pred_masks, gt_masks):
    p = torch.cat([_sod_inputs(a, b)[0] for a, b in zip(pred_masks, gt_masks)], 0)
    g = torch.cat([_sod_inputs(a, b)[1] for a, b in zip(pred_masks, gt_masks)], 0)
    _, part = sod_metrics(p, g)
    return sod_summary_from_partial(part, p.shape[0])
