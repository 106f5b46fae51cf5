"""Batched evaluation pipeline: label maps -> region graphs -> RG-GNN -> RGxKG fusion -> paint + metrics.

This is the composition the reference performs one image at a time across
``extract_embeddings_from_image`` (extract_rg_embeddings.py:249-288), ``predict_single_image``
(test_multimodal.py:83-152) and ``detect_camouflage`` (region_graph/test.py:185-293), run here on a whole
batch without leaving the device: node embeddings go from the GNN straight into the fusion kernels (the
reference round-trips them through host memory, extract_rg_embeddings.py:288 / test_multimodal.py:102).
All buffers are allocated once at capacity; no step synchronises with the host.
"""
from __future__ import annotations

import os

import torch

from . import _lib, metrics
from .region_graph import RegionGraphBatch, build_region_graphs


class EvalPipeline:
    def __init__(self, gnn, detector, kg, label_cap: int = 544, threshold: float = 0.5, label_offset_fix: bool = False,
                 want_pred_mask: bool = False, lane: int = 0, sod_metrics: bool = True, weighted_f: bool = False,
                 table_metrics: bool = True):
        """``kg``: ``[K, kg_dim]`` knowledge-graph rows shared by every image (broadcast once).
        ``lane``: pipelines with different lanes own disjoint scratch and may run concurrently on different streams
        (they share the read-only folded weights).
        ``sod_metrics``: also compute S- / E- / F- / weighted-F-measure (``metrics.SOD_KEYS``; north_star's "S/E/F-measure",
        not in the reference) on the painted mask: ``out["per_image_sod"]`` ``[B,8]``, ``out["partial_sod"]`` ``[528]``.
        ``weighted_f``: also Margolin's weighted F-measure (an exact Euclidean feature transform of every GT mask: several
        times the cost of the other seven; its slot is NaN when off).
        ``table_metrics``: compute the metrics through the region x ground-truth contingency table (``metrics.RegionTable``:
        the pixel pass moves next to the graph build, behind the model only one small block per image remains) instead of
        three pixel passes behind the GNN; same numbers.  Used when the S/E/F metrics are on and neither the painted mask nor
        the weighted F-measure is asked for."""
        self.lane = int(lane)
        self.table_metrics = bool(table_metrics) and os.environ.get("CM_TABLE_METRICS", "1") != "0"
        self._table = None
        self._side2 = None
        self.sod_metrics, self.weighted_f = bool(sod_metrics), bool(weighted_f)
        self.n_segments = 500   # create_region_graph(image, n_segments=500), extract_rg_embeddings.py:138
        self.gnn, self.detector = gnn.eval(), detector.eval()
        kg = kg.reshape(-1, kg.shape[-1])
        _lib.require_cuda(kg, "kg")
        self.kg = _lib.f32c(kg).unsqueeze(0)
        self.label_cap, self.threshold = int(label_cap), float(threshold)
        self.label_offset_fix, self.want_pred_mask = label_offset_fix, want_pred_mask
        self._graph: RegionGraphBatch | None = None
        self._bufs = None
        self._shape = None
        self.total = None      # optional [15] f64 device accumulator of the per-batch metric partial sums
        self.total_sod = None  # optional [528] f64 accumulator of the S/E/F partial sums
        self._side = None      # side stream for the paint + metrics branch
        self.parallel_metrics = os.environ.get("CM_PARALLEL_METRICS", "1") != "0"

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
            "per_image_sod": torch.empty((B, len(metrics.SOD_KEYS)), dtype=torch.float64, device=dev) if self.sod_metrics else None,
            "partial_sod": torch.empty(metrics.SOD_PARTIAL, dtype=torch.float64, device=dev) if self.sod_metrics else None,
        }

    def run(self, images, labels, canny, gt):
        """Device inputs: images [B,H,W,3] f32 (or uint8 = byte/255), labels [B,H,W] i32 (or int16), canny [B,H,W] u8,
        gt [B,H,W] u8.  ``labels=None`` / ``canny=None`` (uint8 images): the front end of the reference's path as shipped runs
        on the device first -- SLIC superpixels (``cm_slic``: n_segments 500, compactness 10, sigma 1,
        extract_rg_embeddings.py:143-144) and the Canny edge map (``cm_canny``, sigma 2, :151-152) -- so only the image
        and the ground truth cross PCIe.  ``canny`` / ``gt`` may instead be bit-packed ``[B, H*W/8]`` uint8 (one bit per pixel,
        ``pack_mask_bits``; a set gt bit means 255): they are unpacked on the device first (``cm_unpack_bits``).
        Returns a dict of device tensors (valid until the next call).  ``out["status"]`` is the graph build's overflow
        word ([1] i32, non-zero = a label >= label_cap was dropped or the node / edge capacity truncated the graph, see
        ``check_status``): the run itself never synchronises, so the caller reads it with its results --
        ``HostPipeline`` copies it back with the logits and raises in ``synchronize()``."""
        from .region_graph import WORKSPACE
        prev_lane, WORKSPACE.lane = WORKSPACE.lane, self.lane
        try:
            return self._run(images, labels, canny, gt)
        finally:
            WORKSPACE.lane = prev_lane

    def _run(self, images, labels, canny, gt):
        lib = _lib.load()
        B, H, W, _ = images.shape
        dev = images.device
        self._alloc(B, H, W, dev)
        if labels is None or canny is None:     # image-only input: device SLIC / Canny
            from .region_graph import canny_edges, slic_segments
            if images.dtype != torch.uint8:
                raise ValueError("the device front end takes uint8 images (the reference hands skimage.slic (image * 255).astype(uint8))")
            if labels is None:
                buf = self._bufs.get("slic_labels")
                if buf is None or buf.shape != (B, H, W):
                    buf = self._bufs["slic_labels"] = torch.empty((B, H, W), dtype=torch.int32, device=dev)
                labels = slic_segments(images, n_segments=self.n_segments, compactness=10.0, sigma=1.0, out=buf)
            if canny is None:
                buf = self._bufs.get("canny_map")
                if buf is None or buf.shape != (B, H, W):
                    buf = self._bufs["canny_map"] = torch.empty((B, H, W), dtype=torch.uint8, device=dev)
                canny = canny_edges(images, sigma=2.0, out=buf)
        if labels is not None and labels.dtype == torch.uint8 and labels.dim() == 2:      # 10-bit label stream (pack_labels10)
            if tuple(labels.shape) != (B, H * W * 10 // 8) or (H * W) % 16:
                raise ValueError(f"packed labels must be [B, H*W*10/8] uint8 with H*W % 16 == 0, got {tuple(labels.shape)}")
            buf = self._bufs.get("labels16")
            if buf is None or buf.shape != (B, H, W):
                buf = self._bufs["labels16"] = torch.empty((B, H, W), dtype=torch.int16, device=dev)
            _lib.check(lib.cm_unpack_labels10(_lib.ptr(labels.contiguous()), B * H * W, _lib.ptr(buf), _lib.stream_ptr(dev)), "cm_unpack_labels10")
            labels = buf
        use_table = (self.table_metrics and self.sod_metrics and not self.weighted_f and not self.want_pred_mask
                     and self.label_cap <= metrics.RT_MAX_LABEL_CAP and 0.0 <= self.threshold < 1.0)
        if canny.dim() == 2 or gt.dim() == 2:      # bit-packed masks: one bit per pixel over PCIe, bytes again here
            for name, t, one in (("canny", canny, 1), ("gt", gt, 255)):
                if t.dim() != 2 or (name == "gt" and use_table):    # the region table reads the packed ground truth as it is
                    continue
                if (H * W) % 8 or tuple(t.shape) != (B, H * W // 8) or t.dtype != torch.uint8:
                    raise ValueError(f"packed {name} must be [B, H*W/8] uint8, got {tuple(t.shape)} {t.dtype}")
                buf = self._bufs.get("unpacked_" + name)
                if buf is None or buf.shape != (B, H, W):
                    buf = self._bufs["unpacked_" + name] = torch.empty((B, H, W), dtype=torch.uint8, device=dev)
                _lib.check(lib.cm_unpack_bits(_lib.ptr(t.contiguous()), t.numel(), one, _lib.ptr(buf), _lib.stream_ptr(dev)), "cm_unpack_bits")
                if name == "canny":
                    canny = buf
                else:
                    gt = buf
        # host-side label maps travel as int16: with uint8 images the int16 kernels read them as they are, otherwise they
        # are widened once on the device
        i16_direct = labels.dtype == torch.int16 and images.dtype == torch.uint8 and W % 8 == 0 and labels.data_ptr() % 16 == 0
        if labels.dtype == torch.int16 and not i16_direct:
            l32 = self._bufs.get("labels32")
            if l32 is None or l32.shape != labels.shape:
                l32 = self._bufs["labels32"] = torch.empty((B, H, W), dtype=torch.int32, device=dev)
            _lib.check(lib.cm_labels_widen_i16(_lib.ptr(labels), labels.numel(), _lib.ptr(l32), _lib.stream_ptr(dev)),
                       "cm_labels_widen_i16")
            labels = l32
        main = torch.cuda.current_stream(dev)
        if use_table:      # labels x gt contingency table: needs nothing from the model, runs next to the graph build
            if self._table is None:
                self._table = metrics.RegionTable(self.label_cap)
            if self._side2 is None or self._side2.device != dev:
                self._side2 = torch.cuda.Stream(dev)
            side2 = self._side2 if self.parallel_metrics else main
            if side2 is not main:
                side2.wait_stream(main)
            with torch.cuda.stream(side2):
                self._table.build(labels, gt)
        g = build_region_graphs(images, labels, canny, label_cap=self.label_cap, want_edge_index=False, out=self._graph)
        self._graph = g
        n_dev = g.node_ptr[B:]
        r = self.gnn.run(g.x, g.csr_row_ptr, g.csr_col, g.edge_w, None, n_dev, g.node_ptr, B,
                         want=("node_emb", "mask_out", "mask_prob"), out=self._bufs["gnn"])
        # paint + metrics need only the GNN's mask probabilities: they run on a side stream next to the fusion (inside a
        # captured step this becomes a parallel branch of the CUDA graph)
        if self._side is None or self._side.device != dev:
            self._side = torch.cuda.Stream(dev)
        side = self._side if self.parallel_metrics else main
        if side is not main:
            side.wait_stream(main)
            if use_table:
                side.wait_stream(self._side2)
        p = _lib.ptr
        b = self._bufs
        if use_table:
            with torch.cuda.stream(side):
                self._table.metrics(r["mask_prob"], g.node_ptr, g.id_of_label if self.label_offset_fix else None, self.threshold,
                                    counts=b["counts"], abs_err=b["abs_err"], per_image=b["per_image"], partial=b["partial"],
                                    per_image_sod=b["per_image_sod"], partial_sod=b["partial_sod"])
                if self.total is not None:
                    self.total.add_(b["partial"])
                    if self.total_sod is not None:
                        self.total_sod.add_(b["partial_sod"])
        else:
            with torch.cuda.stream(side):
                paint = lib.cm_paint_metrics_i16 if labels.dtype == torch.int16 else lib.cm_paint_metrics
                _lib.check(paint(p(labels), p(gt), B, H, W, p(r["mask_prob"]), p(g.node_ptr),
                                                p(g.id_of_label) if self.label_offset_fix else None, self.label_cap,
                                                self.threshold, p(b["pred_mask"]), p(b["counts"]), p(b["abs_err"]),
                                                _lib.stream_ptr(dev)), "cm_paint_metrics")
                _lib.check(lib.cm_metrics_finalize(p(b["counts"]), p(b["abs_err"]), B, H * W, p(b["per_image"]), p(b["partial"]),
                                                   _lib.stream_ptr(dev)), "cm_metrics_finalize")
                if self.sod_metrics:    # on the painted mask, without materialising it
                    metrics.sod_metrics_painted(labels, gt, r["mask_prob"], g.node_ptr, g.id_of_label if self.label_offset_fix else None,
                                                self.label_cap, per_image=b["per_image_sod"], partial=b["partial_sod"],
                                                weighted_f=self.weighted_f)
                if self.total is not None:
                    self.total.add_(b["partial"])
                    if self.total_sod is not None and self.sod_metrics:
                        self.total_sod.add_(b["partial_sod"])
        f = self.detector.forward_ragged(r["node_emb"], g.node_ptr, B, self.kg, return_attention=False,
                                         out=self._bufs["fusion"])
        self._bufs["fusion"] = f
        if side is not main:
            main.wait_stream(side)
        return {"graph": g, "status": g.status, "node_emb": r["node_emb"], "node_mask_out": r["mask_out"], "mask_prob": r["mask_prob"],
                "mask_out": f["mask_out"], "instance_out": f["instance_out"], "edge_out": f["edge_out"],
                "score_out": f["score_out"], "fused": f["fused"], "counts": b["counts"], "abs_err": b["abs_err"],
                "per_image": b["per_image"], "partial": b["partial"], "pred_mask": b["pred_mask"],
                "per_image_sod": b["per_image_sod"], "partial_sod": b["partial_sod"]}


    def capture(self, images, labels, canny, gt):
        """Record one ``run`` over the given (static) device buffers into a CUDA graph: the ~45 kernel launches of a
        step become one ``cudaGraphLaunch``.  Returns ``(graph, out)``; ``graph.replay()`` recomputes ``out`` in place
        from whatever the input buffers hold at that time."""
        _lib.timing_enable(False)
        side = torch.cuda.Stream(images.device)
        side.wait_stream(torch.cuda.current_stream(images.device))
        with torch.cuda.stream(side):          # warm-up outside capture: buffers, tensor maps, function attributes
            for _ in range(2):
                self.run(images, labels, canny, gt)
        torch.cuda.current_stream(images.device).wait_stream(side)
        torch.cuda.synchronize(images.device)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            out = self.run(images, labels, canny, gt)
        return graph, out


def check_status(status: int, what: str = "region-graph build") -> None:
    """Raise when the overflow word of a graph build is non-zero: the logits / metrics of that batch are not valid."""
    status = int(status)
    if status:
        names = ", ".join(n for b, n in ((_lib.STATUS_LABEL_OVERFLOW, "label >= label_cap (region dropped)"),
                                         (_lib.STATUS_NODE_OVERFLOW, "nodes > node_cap"), (_lib.STATUS_EDGE_OVERFLOW, "edges > edge_cap"))
                          if status & b)
        raise _lib.CmError(f"{what} overflow (status {status}): {names}")


def pack_mask_bits(mask: torch.Tensor) -> torch.Tensor:
    """``[B,H,W]`` mask (any non-zero = set) -> ``[B, H*W/8]`` uint8, bit i of byte j = pixel 8 j + i (little-endian bit
    order, what ``cm_unpack_bits`` reads).  Host-side helper for binary masks (a Canny map, a {0,255} ground truth)."""
    import numpy as np
    m = mask.detach().cpu().numpy() != 0
    return torch.from_numpy(np.packbits(m.reshape(m.shape[0], -1), axis=1, bitorder="little"))


def pack_labels10(labels: torch.Tensor) -> torch.Tensor:
    """``[B,H,W]`` integer label map with values in [0, 1024) -> ``[B, H*W*10/8]`` uint8: a little-endian bit stream, label i
    in bits [10 i, 10 i + 10) (what ``cm_unpack_labels10`` reads).  Host-side helper: 1.25 instead of 2 bytes per pixel."""
    import numpy as np
    lab = labels.detach().cpu().numpy()
    B = lab.shape[0]
    if lab.size and (lab.min() < 0 or lab.max() > 1023):
        raise ValueError("pack_labels10: labels must lie in [0, 1024)")
    if (lab.size // max(B, 1)) % 16:
        raise ValueError("pack_labels10: H*W must be a multiple of 16")
    q = lab.reshape(B, -1, 4).astype(np.uint16)
    l0, l1, l2, l3 = q[..., 0], q[..., 1], q[..., 2], q[..., 3]
    out = np.empty((B, q.shape[1], 5), np.uint8)
    out[..., 0] = l0 & 255
    out[..., 1] = (l0 >> 8) | ((l1 & 63) << 2)
    out[..., 2] = (l1 >> 6) | ((l2 & 15) << 4)
    out[..., 3] = (l2 >> 4) | ((l3 & 3) << 6)
    out[..., 4] = l3 >> 2
    return torch.from_numpy(out.reshape(B, -1))


class BatchBuffer:
    """The four inputs of one batch as views of ONE flat byte buffer (pinned host memory or device memory), each at a
    256-byte aligned offset: a host batch laid out like this goes to the device as a single copy (54.7 GB/s for
    14.7 MB on this machine's PCIe link against 50-54 GB/s for four separate 2-6 MB copies).  ``packed_masks``: the
    Canny map and the ground truth travel one bit per pixel (``[B, H*W/8]`` uint8; 7 -> 5.25 bytes per pixel with uint8
    images + int16 labels) -- for binary masks only: a set gt bit is read as 255."""

    def __init__(self, B, H, W, image_dtype=torch.float32, label_dtype=torch.int32, device=None, packed_masks=False, image_only=False,
                 packed_labels=False):
        """``image_only``: no label / Canny views -- the device front end (cm_slic, cm_canny) derives both from the image,
        so a batch is uint8 RGB + the ground truth (3.125 bytes per pixel with a bit-packed mask).
        ``packed_labels``: the label map travels at 10 bits per pixel (``[B, H*W*10/8]`` uint8, ``pack_labels10``; labels
        below 1024): 4.5 bytes per pixel with uint8 images and bit-packed masks."""
        self.packed_masks, self.image_only = bool(packed_masks), bool(image_only)
        self.packed_labels = bool(packed_labels) and not self.image_only
        if self.packed_labels and (H * W) % 16:
            raise ValueError("packed labels need H*W % 16 == 0")
        mshape = (B, H * W // 8) if packed_masks else (B, H, W)
        if packed_masks and (H * W) % 8:
            raise ValueError("packed masks need H*W % 8 == 0")
        shapes = [("images", (B, H, W, 3), image_dtype)]
        if not image_only:
            shapes += [("labels", (B, H * W * 10 // 8), torch.uint8) if self.packed_labels else ("labels", (B, H, W), label_dtype),
                       ("canny", mshape, torch.uint8)]
        shapes.append(("gt", mshape, torch.uint8))
        offs, total = [], 0
        for _, shp, dt in shapes:
            offs.append(total)
            n = 1
            for d in shp:
                n *= d
            total += (n * torch.empty((), dtype=dt).element_size() + 255) // 256 * 256
        self.flat = (torch.empty(total, dtype=torch.uint8, pin_memory=True) if device is None
                     else torch.empty(total, dtype=torch.uint8, device=device))
        self.images = self.labels = self.canny = self.gt = None
        self.payload_bytes = 0
        for (name, shp, dt), o in zip(shapes, offs):
            n = 1
            for d in shp:
                n *= d
            v = self.flat[o:o + n * torch.empty((), dtype=dt).element_size()].view(dt).view(shp)
            setattr(self, name, v)
            self.payload_bytes += v.numel() * v.element_size()

    def fill(self, images, labels, canny, gt):
        """Copy a batch in; with ``packed_masks`` byte masks ``[B,H,W]`` are bit-packed on the way (host work); with
        ``image_only`` the label and Canny maps are ignored."""
        self.images.copy_(images)
        if not self.image_only:
            self.labels.copy_(pack_labels10(labels) if self.packed_labels and labels.dim() == 3 else labels)
            self.canny.copy_(pack_mask_bits(canny) if self.packed_masks and canny.dim() == 3 else canny)
        self.gt.copy_(pack_mask_bits(gt) if self.packed_masks and gt.dim() == 3 else gt)
        return self


class HostStager:
    """Pinned host buffers + device buffers for one batch: the host-facing entry of the pipeline."""

    def __init__(self, B, H, W, device, image_dtype=torch.float32, label_dtype=torch.int32, packed_masks=False, image_only=False,
                 packed_labels=False):
        self.dev = device
        pin = dict(pin_memory=True)
        self.h_batch = BatchBuffer(B, H, W, image_dtype, label_dtype, packed_masks=packed_masks, image_only=image_only, packed_labels=packed_labels)
        self.d_batch = BatchBuffer(B, H, W, image_dtype, label_dtype, device=device, packed_masks=packed_masks, image_only=image_only,
                                   packed_labels=packed_labels)
        self.h_images, self.h_labels, self.h_canny, self.h_gt = (self.h_batch.images, self.h_batch.labels, self.h_batch.canny,
                                                                 self.h_batch.gt)
        self.d_images, self.d_labels, self.d_canny, self.d_gt = (self.d_batch.images, self.d_batch.labels, self.d_batch.canny,
                                                                 self.d_batch.gt)
        self.h_mask_out = torch.empty((B, 2), dtype=torch.float32, **pin)
        self.h_per_image = torch.empty((B, 7), dtype=torch.float64, **pin)
        self.h_status = torch.zeros(1, dtype=torch.int32, **pin)
        self.h_per_image_sod = torch.empty((B, len(metrics.SOD_KEYS)), dtype=torch.float64, **pin)
        self.h2d_bytes = self.h_batch.payload_bytes
        self.d2h_bytes = sum(t.numel() * t.element_size() for t in (self.h_mask_out, self.h_per_image, self.h_status, self.h_per_image_sod))

    def fill(self, batch):
        self.h_batch.fill(batch.images, batch.labels, batch.canny, batch.gt)

    def step(self, pipe: EvalPipeline):
        """One end-to-end step from pinned host memory: H2D (one copy), the whole path, D2H of logits + metrics."""
        self.d_batch.flat.copy_(self.h_batch.flat, non_blocking=True)
        out = pipe.run(self.d_images, self.d_labels, self.d_canny, self.d_gt)
        self.h_mask_out.copy_(out["mask_out"], non_blocking=True)
        self.h_per_image.copy_(out["per_image"], non_blocking=True)
        self.h_status.copy_(out["status"], non_blocking=True)
        if out["per_image_sod"] is not None:
            self.h_per_image_sod.copy_(out["per_image_sod"], non_blocking=True)
        return out

    def check(self):
        """After the step's stream work has completed: raise if the graph build overflowed."""
        check_status(self.h_status.item())


class HostPipeline:
    """Host-facing entry of the evaluation path: batches arrive in pinned host memory and each goes through one
    *slot* -- its own stream, device input buffers, captured step and pinned result buffers -- as H2D copy ->
    captured step -> D2H of the logits + per-image metrics.  Slots are rotated, so the copies of one batch overlap
    the kernels of the others; slots whose pipelines sit on different lanes (``pipes`` list) also overlap their
    kernels, slots sharing a pipeline are serialised through ``pipe_done`` events (they share its buffers)."""

    def __init__(self, pipe, B, H, W, device, slots: int = 2, use_graph: bool = True, image_dtype=torch.float32,
                 label_dtype=torch.int32, packed_masks: bool = False, image_only: bool = False, ingest: str | None = None,
                 packed_labels: bool = False):
        """``ingest``: how a ``BatchBuffer`` crosses PCIe -- ``"ce"`` (cudaMemcpyAsync: a copy engine; default, env
        ``CM_E2E_INGEST``), ``"sm"`` (``cm_ingest_copy``: a 64-block kernel reading the pinned buffer) or ``"auto"`` (both are
        timed at construction under the load of the slots' own captured steps, the faster one is kept).  The copy engine
        is the faster path (54 against 47 GB/s on an idle link here, 146 k against 97 k img/s end to end); the kernel form
        exists for callers that want the transfer inside a captured graph.  Note for the producer of the batches: a pinned
        buffer the CPU has JUST written is still dirty in its caches and is read by the device at 10-30 GB/s instead of 54
        (``bench.py``: ``h2d_probe_per_buffer_gbs``) -- write staging buffers with streaming stores or a few batches ahead."""
        pipes = list(pipe) if isinstance(pipe, (list, tuple)) else [pipe]
        self.pipes, self.dev = pipes, device
        # One stream carries every host->device copy, in submission order: each batch crosses PCIe at the link's full rate and
        # the batch that is needed first arrives first.  (With the copy on the slot's own stream the copies of several slots
        # share the link and all of them finish late.)  CM_E2E_COPY_STREAM=0: the copy on the slot's stream.
        self.copy_stream = torch.cuda.Stream(device) if os.environ.get("CM_E2E_COPY_STREAM", "1") != "0" else None
        self.pipe_done = [torch.cuda.Event() for _ in pipes]
        self.slots = []
        for k in range(slots):
            st = HostStager(B, H, W, device, image_dtype, label_dtype, packed_masks=packed_masks, image_only=image_only, packed_labels=packed_labels)
            st.pipe_idx = k % len(pipes)
            st.stream = torch.cuda.Stream(device)
            st.d_batch.flat.zero_()
            if st.d_labels is not None:
                st.d_labels.fill_(1)
            st.ev_done = torch.cuda.Event()
            st.ev_copied, st.ev_consumed = torch.cuda.Event(), torch.cuda.Event()
            st.graph, st.out = pipes[st.pipe_idx].capture(st.d_images, st.d_labels, st.d_canny, st.d_gt) if use_graph else (None, None)
            st.ev_done.record(torch.cuda.current_stream(device))
            st.ev_consumed.record(torch.cuda.current_stream(device))
            self.slots.append(st)
        for ev in self.pipe_done:
            ev.record(torch.cuda.current_stream(device))
        self.h2d_bytes, self.d2h_bytes = self.slots[0].h2d_bytes, self.slots[0].d2h_bytes
        ingest = ingest or os.environ.get("CM_E2E_INGEST", "ce")
        if ingest not in ("ce", "sm", "auto"):
            raise ValueError("ingest must be 'ce', 'sm' or 'auto'")
        flat = self.slots[0].d_batch.flat
        sm_ok = flat.numel() % 16 == 0 and flat.data_ptr() % 16 == 0 and self.slots[0].h_batch.flat.data_ptr() % 16 == 0
        self.ingest_rates = None
        if ingest == "auto":
            ingest = self._calibrate() if (sm_ok and use_graph) else "ce"
        self.ingest = ingest if sm_ok else "ce"

    def _stage(self, stream, h_flat, d_flat, mode):
        """Queue the host -> device transfer of one flat batch on ``stream`` (current stream = ``stream``)."""
        if mode == "sm":
            lib = _lib.load()
            _lib.check(lib.cm_ingest_copy(h_flat.data_ptr(), _lib.ptr(d_flat), d_flat.numel(), 64, _lib.stream_ptr(self.dev)), "cm_ingest_copy")
        else:
            d_flat.copy_(h_flat, non_blocking=True)

    def _calibrate(self, rounds: int = 3) -> str:
        """Time ``rounds * slots`` end-to-end submissions with each transfer mode.  The transfers land in a scratch buffer, so
        the slots' device inputs (valid data from the capture) are not disturbed while their captured steps run as the load."""
        import time
        scratch = torch.empty_like(self.slots[0].d_batch.flat)
        rates = {}
        for mode in ("ce", "sm", "ce", "sm"):
            torch.cuda.synchronize(self.dev)
            t0 = time.perf_counter()
            for i in range(rounds * len(self.slots)):
                st = self.slots[i % len(self.slots)]
                cs = self.copy_stream if self.copy_stream is not None else st.stream
                with torch.cuda.stream(cs):
                    cs.wait_event(st.ev_consumed)
                    self._stage(cs, st.h_batch.flat, scratch, mode)
                    st.ev_copied.record(cs)
                with torch.cuda.stream(st.stream):
                    st.stream.wait_event(st.ev_copied)
                    st.stream.wait_event(self.pipe_done[st.pipe_idx])
                    st.graph.replay()
                    self.pipe_done[st.pipe_idx].record(st.stream)
                    st.ev_consumed.record(st.stream)
                    st.ev_done.record(st.stream)
            torch.cuda.synchronize(self.dev)
            rates[mode] = max(rates.get(mode, 0.0), rounds * len(self.slots) / (time.perf_counter() - t0))   # steps / s, best of two
        self.ingest_rates = rates
        return "sm" if rates["sm"] > 1.10 * rates["ce"] else "ce"

    def host_batch(self) -> BatchBuffer:
        """A pinned host batch in the layout of the device slots: fill its ``images / labels / canny / gt`` views (e.g.
        decode straight into them) and pass it to ``submit`` -- it then crosses PCIe as ONE copy."""
        st = self.slots[0]
        return BatchBuffer(*st.h_images.shape[:3], st.h_images.dtype, st.h_labels.dtype if st.h_labels is not None else torch.int32,
                           packed_masks=st.h_batch.packed_masks, image_only=st.h_batch.image_only, packed_labels=st.h_batch.packed_labels)

    def submit(self, i, h_images, h_labels=None, h_canny=None, h_gt=None):
        """Queue batch ``i``: a ``BatchBuffer`` from ``host_batch()`` (one H2D copy) or four pinned host tensors (four
        copies); returns its slot.  The slot's host results are valid once ``slot.ev_done`` has completed and until
        the slot is reused ``len(slots)`` submissions later."""
        st = self.slots[i % len(self.slots)]
        pipe = self.pipes[st.pipe_idx]
        cs = self.copy_stream if self.copy_stream is not None else st.stream
        with torch.cuda.stream(cs):
            if cs is not st.stream:
                cs.wait_event(st.ev_consumed)      # the slot's previous step has read its device inputs
            if isinstance(h_images, BatchBuffer):
                self._stage(cs, h_images.flat, st.d_batch.flat, self.ingest if h_images.flat.data_ptr() % 16 == 0 else "ce")
            else:
                st.d_images.copy_(h_images, non_blocking=True)
                if st.d_labels is not None:
                    st.d_labels.copy_(h_labels, non_blocking=True)
                    st.d_canny.copy_(h_canny, non_blocking=True)
                st.d_gt.copy_(h_gt, non_blocking=True)
            if cs is not st.stream:
                st.ev_copied.record(cs)
        with torch.cuda.stream(st.stream):
            if cs is not st.stream:
                st.stream.wait_event(st.ev_copied)
            st.stream.wait_event(self.pipe_done[st.pipe_idx])     # another slot may be using this pipeline's buffers
            if st.graph is not None:
                st.graph.replay()
                out = st.out
            else:
                out = pipe.run(st.d_images, st.d_labels, st.d_canny, st.d_gt)
            st.h_mask_out.copy_(out["mask_out"], non_blocking=True)
            st.h_per_image.copy_(out["per_image"], non_blocking=True)
            st.h_status.copy_(out["status"], non_blocking=True)
            if out["per_image_sod"] is not None:
                st.h_per_image_sod.copy_(out["per_image_sod"], non_blocking=True)
            self.pipe_done[st.pipe_idx].record(st.stream)
            st.ev_consumed.record(st.stream)
            st.ev_done.record(st.stream)
        return st

    def synchronize(self):
        """Wait for every slot, then raise if any batch's graph build overflowed (its logits / metrics are invalid)."""
        for st in self.slots:
            st.ev_done.synchronize()
        for st in self.slots:
            st.check()


def bind_host_to_gpu(device_index: int) -> list[int]:
    """Pin this process (and the pinned staging buffers it first-touches afterwards) to the CPU cores NVML reports as
    local to the GPU, so host->device copies do not cross the socket interconnect.  Returns the cores bound to
    (empty when NVML or the affinity call is unavailable -- the process then keeps its inherited affinity)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        handle = pynvml.nvmlDeviceGetHandleByIndex(_physical_index(device_index))
        words = pynvml.nvmlDeviceGetCpuAffinity(handle, (os.cpu_count() + 63) // 64)
        cores = [64 * i + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1]
        allowed = sorted(set(cores) & os.sched_getaffinity(0))
        if allowed:
            os.sched_setaffinity(0, allowed)
        return allowed
    except Exception:   # noqa: BLE001 -- placement is an optimisation, never a requirement
        return []


def _physical_index(device_index: int) -> int:
    vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
    ids = [v for v in vis.split(",") if v.strip()]
    if ids and device_index < len(ids) and ids[device_index].strip().isdigit():
        return int(ids[device_index])
    return device_index


def allreduce_partials(partial: torch.Tensor, confusion: torch.Tensor | None = None, partial_sod: torch.Tensor | None = None):
    """Sum the metric partial sums (and image-level confusion counts, and the S/E/F partial vector) over ranks: the
    only data-path collective of the sharded evaluation.  No-op without an initialised process group.
    Returns ``(partial, confusion)`` -- plus ``partial_sod`` when one was given."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        if partial_sod is not None:      # one collective for both float64 vectors
            both = torch.cat([partial, partial_sod])
            dist.all_reduce(both, op=dist.ReduceOp.SUM)
            partial.copy_(both[:partial.numel()])
            partial_sod.copy_(both[partial.numel():])
        else:
            dist.all_reduce(partial, op=dist.ReduceOp.SUM)
        if confusion is not None:
            dist.all_reduce(confusion, op=dist.ReduceOp.SUM)
    return (partial, confusion) if partial_sod is None else (partial, confusion, partial_sod)


def shard_indices(n_items: int, rank: int, world: int):
    """Image index -> rank: ``i mod world``."""
    return list(range(rank, n_items, world))


def summarize(partial) -> dict:
    return metrics.summary_from_partial(partial)
