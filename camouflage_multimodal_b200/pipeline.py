"""Batched evaluation pipeline: label maps -> region graphs -> RG-GNN -> RGxKG fusion -> paint + metrics.

This is the composition the reference performs one image at a time across
``extract_embeddings_from_image`` (extract_rg_embeddings.py:249-288), ``predict_single_image``
(test_multimodal.py:83-152) and ``detect_camouflage`` (region_graph/test.py:185-293), run here on a whole
batch without leaving the device: node embeddings go from the GNN straight into the fusion kernels (the
reference round-trips them through host memory, extract_rg_embeddings.py:288 / test_multimodal.py:102).
All buffers are allocated once at capacity; no step synchronises with the host.
"""
from __future__ import annotations

import torch

from . import _lib, metrics
from .region_graph import RegionGraphBatch, build_region_graphs


class EvalPipeline:
    def __init__(self, gnn, detector, kg, label_cap: int = 544, threshold: float = 0.5, label_offset_fix: bool = False,
                 want_pred_mask: bool = False):
        """``kg``: ``[K, kg_dim]`` knowledge-graph rows shared by every image (broadcast once)."""
        self.gnn, self.detector = gnn.eval(), detector.eval()
        kg = kg.reshape(-1, kg.shape[-1])
        _lib.require_cuda(kg, "kg")
        self.kg = _lib.f32c(kg).unsqueeze(0)
        self.label_cap, self.threshold = int(label_cap), float(threshold)
        self.label_offset_fix, self.want_pred_mask = label_offset_fix, want_pred_mask
        self._graph: RegionGraphBatch | None = None
        self._bufs = None
        self._shape = None

    def _alloc(self, B, H, W, dev):
        if self._shape == (B, H, W, str(dev)):
            return
        self._graph = None
        self._shape = (B, H, W, str(dev))
        f32 = dict(dtype=torch.float32, device=dev)
        node_cap = B * min(self.label_cap, H * W)
        h, c = self.gnn.hidden_channels, self.gnn.num_classes
        self._bufs = {
            "gnn": {"node_emb": torch.empty((node_cap, h), **f32), "mask_out": torch.empty((node_cap, c), **f32),
                    "mask_prob": torch.empty(node_cap, **f32)},
            "fusion": {},
            "counts": torch.empty((B, 4), dtype=torch.int64, device=dev),
            "abs_err": torch.empty(B, dtype=torch.float64, device=dev),
            "per_image": torch.empty((B, 7), dtype=torch.float64, device=dev),
            "partial": torch.empty(15, dtype=torch.float64, device=dev),
            "pred_mask": torch.empty((B, H, W), **f32) if self.want_pred_mask else None,
        }

    def run(self, images, labels, canny, gt):
        """Device inputs: images [B,H,W,3] f32, labels [B,H,W] i32, canny [B,H,W] u8, gt [B,H,W] u8.
        Returns a dict of device tensors (valid until the next call)."""
        lib = _lib.load()
        B, H, W, _ = images.shape
        dev = images.device
        self._alloc(B, H, W, dev)
        g = build_region_graphs(images, labels, canny, label_cap=self.label_cap, want_edge_index=False, out=self._graph)
        self._graph = g
        n_dev = g.node_ptr[B:]
        r = self.gnn.run(g.x, g.csr_row_ptr, g.csr_col, g.edge_w, None, n_dev, g.node_ptr, B,
                         want=("node_emb", "mask_out", "mask_prob"), out=self._bufs["gnn"])
        f = self.detector.forward_ragged(r["node_emb"], g.node_ptr, B, self.kg, return_attention=False,
                                         out=self._bufs["fusion"])
        self._bufs["fusion"] = f
        p = _lib.ptr
        b = self._bufs
        _lib.check(lib.cm_paint_metrics(p(labels), p(gt), B, H, W, p(r["mask_prob"]), p(g.node_ptr),
                                        p(g.id_of_label) if self.label_offset_fix else None, self.label_cap,
                                        self.threshold, p(b["pred_mask"]), p(b["counts"]), p(b["abs_err"]),
                                        _lib.stream_ptr(dev)), "cm_paint_metrics")
        _lib.check(lib.cm_metrics_finalize(p(b["counts"]), p(b["abs_err"]), B, H * W, p(b["per_image"]), p(b["partial"]),
                                           _lib.stream_ptr(dev)), "cm_metrics_finalize")
        return {"graph": g, "node_emb": r["node_emb"], "node_mask_out": r["mask_out"], "mask_prob": r["mask_prob"],
                "mask_out": f["mask_out"], "instance_out": f["instance_out"], "edge_out": f["edge_out"],
                "score_out": f["score_out"], "fused": f["fused"], "counts": b["counts"], "abs_err": b["abs_err"],
                "per_image": b["per_image"], "partial": b["partial"], "pred_mask": b["pred_mask"]}


class HostStager:
    """Pinned host buffers + device buffers for one batch: the host-facing entry of the pipeline."""

    def __init__(self, B, H, W, device):
        self.dev = device
        pin = dict(pin_memory=True)
        self.h_images = torch.empty((B, H, W, 3), dtype=torch.float32, **pin)
        self.h_labels = torch.empty((B, H, W), dtype=torch.int32, **pin)
        self.h_canny = torch.empty((B, H, W), dtype=torch.uint8, **pin)
        self.h_gt = torch.empty((B, H, W), dtype=torch.uint8, **pin)
        self.d_images = torch.empty((B, H, W, 3), dtype=torch.float32, device=device)
        self.d_labels = torch.empty((B, H, W), dtype=torch.int32, device=device)
        self.d_canny = torch.empty((B, H, W), dtype=torch.uint8, device=device)
        self.d_gt = torch.empty((B, H, W), dtype=torch.uint8, device=device)
        self.h_mask_out = torch.empty((B, 2), dtype=torch.float32, **pin)
        self.h_per_image = torch.empty((B, 7), dtype=torch.float64, **pin)
        self.h2d_bytes = sum(t.numel() * t.element_size() for t in (self.h_images, self.h_labels, self.h_canny, self.h_gt))
        self.d2h_bytes = sum(t.numel() * t.element_size() for t in (self.h_mask_out, self.h_per_image))

    def fill(self, batch):
        self.h_images.copy_(batch.images)
        self.h_labels.copy_(batch.labels)
        self.h_canny.copy_(batch.canny)
        self.h_gt.copy_(batch.gt)

    def step(self, pipe: EvalPipeline):
        """One end-to-end step from pinned host memory: H2D, the whole path, D2H of logits + metrics."""
        self.d_images.copy_(self.h_images, non_blocking=True)
        self.d_labels.copy_(self.h_labels, non_blocking=True)
        self.d_canny.copy_(self.h_canny, non_blocking=True)
        self.d_gt.copy_(self.h_gt, non_blocking=True)
        out = pipe.run(self.d_images, self.d_labels, self.d_canny, self.d_gt)
        self.h_mask_out.copy_(out["mask_out"], non_blocking=True)
        self.h_per_image.copy_(out["per_image"], non_blocking=True)
        return out


def allreduce_partials(partial: torch.Tensor, confusion: torch.Tensor | None = None):
    """Sum the metric partial sums (and image-level confusion counts) over ranks: the only data-path collective
    of the sharded evaluation.  No-op without an initialised process group."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(partial, op=dist.ReduceOp.SUM)
        if confusion is not None:
            dist.all_reduce(confusion, op=dist.ReduceOp.SUM)
    return partial, confusion


def shard_indices(n_items: int, rank: int, world: int):
    """Image index -> rank: ``i mod world``."""
    return list(range(rank, n_items, world))


def summarize(partial) -> dict:
    return metrics.summary_from_partial(partial)
