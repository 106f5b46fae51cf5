"""Region-graph build and RegionGraphGNN on the B200 kernels.

Mirrors ``models/region_graph/extract_rg_embeddings.py`` of the reference: ``create_region_graph`` (:138-246)
from the label map onward, and ``RegionGraphGNN`` (:27-135) with the same constructor, methods, return
arities and PyG state-dict key names (``conv1.lin.weight``, ``conv1.att_src`` ...).  The two steps before the hot
path -- SLIC (``:143-144``) and Canny (8f N2, ``:152``) -- run on the device too (``slic_segments`` /
``cm_slic``, ``canny_edges`` / ``cm_canny``) when the caller does not pass a label map / edge map.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np
import torch
import torch.nn as nn

from . import _lib
from .data import Batch, Data

NORM_SIDE = 256.0   # hard-coded by the reference (extract_rg_embeddings.py:172-174)


class _Workspace:
    """Grow-only byte buffers per (device, tag, lane): scratch handed to the C ABI.  ``lane`` separates
    pipelines that run concurrently on different streams (pipeline.EvalPipeline(lane=...))."""

    def __init__(self):
        self._bufs = {}
        self.lane = 0

    def get(self, device, tag: str, nbytes: int) -> torch.Tensor:
        key = (str(device), tag, self.lane)
        buf = self._bufs.get(key)
        if buf is None or buf.numel() < nbytes:
            buf = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)
            self._bufs[key] = buf
        return buf


WORKSPACE = _Workspace()


@dataclass
class RegionGraphBatch:
    """Device-resident result of one batched graph build (capacity-sized buffers + device-side counts)."""
    x: torch.Tensor             # [node_cap, 15] f32
    edge_w: torch.Tensor        # [edge_cap] f32, one per directed CSR entry
    csr_row_ptr: torch.Tensor   # [node_cap+1] i32
    csr_col: torch.Tensor       # [edge_cap] i32 (global node ids)
    edge_index: torch.Tensor | None   # [2, edge_cap] i64, canonical order
    node_ptr: torch.Tensor      # [B+1] i32
    edge_ptr: torch.Tensor      # [B+1] i32
    id_of_label: torch.Tensor   # [B, label_cap] i32
    node_label: torch.Tensor    # [node_cap] i32
    node_bbox: torch.Tensor     # [node_cap, 4] i32
    batch_vec: torch.Tensor | None    # [node_cap] i64
    status: torch.Tensor        # [1] i32 bit mask
    batch: int
    label_cap: int
    node_cap: int
    edge_cap: int

    def counts(self):
        """(total nodes, total edges): one device->host read."""
        s, n, e = torch.cat([self.status.view(1), self.node_ptr[-1:], self.edge_ptr[-1:]]).tolist()
        if s:
            raise _lib.CmError(
                f"region-graph build overflow (status {s}): "
                + ", ".join(n for b, n in ((1, "label >= label_cap"), (2, "nodes > node_cap"), (4, "edges > edge_cap")) if s & b))
        return int(n), int(e)


def build_region_graphs(images, labels, canny, label_cap: int | None = None, edge_cap: int | None = None,
                        want_edge_index: bool = True, local_ids: bool = False, out: RegionGraphBatch | None = None,
                        norm_side: float = NORM_SIDE, feature_kernel: str = "auto") -> RegionGraphBatch:
    """A1-A4 for a batch: images ``[B,H,W,3]`` f32 in [0,1] or uint8 (value = byte / 255, integer-exact feature
    kernel), labels ``[B,H,W]`` i32 (or int16: read as they are by the ``*_i16`` kernels when the images are uint8 and
    ``W % 8 == 0``, widened otherwise), canny ``[B,H,W]`` u8, all CUDA.

    ``label_cap`` bounds the label values (labels >= label_cap are reported through ``status``); the default
    reads ``labels.max()+1`` back from the device (one sync) -- pass it to stay asynchronous.  ``edge_cap``
    defaults to 16 directed edges per possible node.  No host synchronisation happens when ``label_cap`` is given.
    ``feature_kernel``: ``"auto"`` = pixel-parallel kernel with per-label shared-memory tables (falls back by itself
    when the table does not fit), ``"bbox"`` = one warp per region over its bounding box.  uint8 images default to the
    region-parallel integer kernel; ``"tile"`` selects their pixel-parallel variant.
    """
    lib = _lib.load()
    for t, n in ((images, "images"), (labels, "labels"), (canny, "canny")):
        _lib.require_cuda(t, n)
    if images.dim() != 4 or images.shape[-1] != 3 or images.dtype not in (torch.float32, torch.uint8):
        raise ValueError(f"images must be [B,H,W,3] float32 or uint8, got {tuple(images.shape)} {images.dtype}")
    B, H, W, _ = images.shape
    if tuple(labels.shape) != (B, H, W) or labels.dtype not in (torch.int32, torch.int16):
        raise ValueError(f"labels must be [B,H,W] int32 (or int16), got {tuple(labels.shape)} {labels.dtype}")
    if tuple(canny.shape) != (B, H, W) or canny.dtype != torch.uint8:
        raise ValueError(f"canny must be [B,H,W] uint8, got {tuple(canny.shape)} {canny.dtype}")
    images, labels, canny = images.contiguous(), labels.contiguous(), canny.contiguous()
    dev = images.device
    if label_cap is None:
        label_cap = max(int(labels.max().item()) + 1, 1)
    if out is not None and (out.batch, out.label_cap) == (B, label_cap):
        g = out
        g.status.zero_()
    else:
        node_cap = B * min(label_cap, H * W)
        ecap = edge_cap if edge_cap is not None else 16 * node_cap
        i32 = dict(dtype=torch.int32, device=dev)
        g = RegionGraphBatch(
            x=torch.empty((node_cap, _lib.NODE_FEATURES), dtype=torch.float32, device=dev),
            edge_w=torch.empty(ecap, dtype=torch.float32, device=dev),
            csr_row_ptr=torch.empty(node_cap + 1, **i32), csr_col=torch.empty(ecap, **i32),
            edge_index=torch.empty((2, ecap), dtype=torch.int64, device=dev) if want_edge_index else None,
            node_ptr=torch.empty(B + 1, **i32), edge_ptr=torch.empty(B + 1, **i32),
            id_of_label=torch.empty((B, label_cap), **i32), node_label=torch.empty(node_cap, **i32),
            node_bbox=torch.empty((node_cap, 4), **i32),
            batch_vec=torch.empty(node_cap, dtype=torch.int64, device=dev),
            status=torch.zeros(1, **i32), batch=B, label_cap=label_cap, node_cap=node_cap, edge_cap=ecap)
    st = _lib.stream_ptr(dev)
    tile_cap = 0 if feature_kernel == "bbox" else label_cap
    ws_bytes = lib.cm_rag_workspace_bytes(B, label_cap)
    ws = WORKSPACE.get(dev, "rag", ws_bytes)
    p = _lib.ptr
    # int16 label maps are read as they are (no widened copy) when the int16 kernels apply: uint8 images through the
    # region feature kernel, W % 8 == 0; any other combination widens first
    i16 = labels.dtype == torch.int16
    if i16 and not (images.dtype == torch.uint8 and feature_kernel != "tile" and W % 8 == 0 and labels.data_ptr() % 16 == 0):
        labels, i16 = labels.to(torch.int32), False
    rag_build = lib.cm_rag_build_i16 if i16 else lib.cm_rag_build
    _lib.check(rag_build(p(labels), B, H, W, label_cap, p(ws), ws_bytes, p(g.id_of_label), p(g.node_ptr),
                                p(g.edge_ptr), p(g.node_label), p(g.node_bbox), p(g.batch_vec), p(g.csr_row_ptr),
                                p(g.csr_col), p(g.edge_index), 1 if local_ids else 0, g.node_cap, g.edge_cap,
                                p(g.status), st), "cm_rag_build")
    if images.dtype == torch.uint8:
        feat = lib.cm_node_features_u8_i16 if i16 else lib.cm_node_features_u8
        _lib.check(feat(p(images), p(labels), p(canny), B, H, W, label_cap, p(g.node_ptr), p(g.node_label),
                                           None if feature_kernel == "tile" else p(g.node_bbox), float(norm_side), p(g.x),
                                           g.node_cap, st), "cm_node_features_u8")
    else:
        _lib.check(lib.cm_node_features(p(images), p(labels), p(canny), B, H, W, tile_cap, p(g.node_ptr), p(g.node_label),
                                        p(g.node_bbox), float(norm_side), p(g.x), g.node_cap, st), "cm_node_features")
    n_dev = g.node_ptr[B:]
    _lib.check(lib.cm_edge_weights(p(g.x), p(g.csr_row_ptr), p(g.csr_col), p(n_dev), g.node_cap, p(g.edge_w), st),
               "cm_edge_weights")
    return g


def canny_edges(images, sigma: float = 2.0, low_threshold: float = 0.1, high_threshold: float = 0.2, out=None):
    """``skimage.feature.canny(gray, sigma)`` of every image (extract_rg_embeddings.py:151-152) on the device:
    ``images`` ``[B,H,W,3]`` float32 in [0,1] or uint8 (value = byte/255) -> ``[B,H,W]`` uint8 edge map (0/1).
    Thresholds are skimage's defaults for float images (0.1 / 0.2 of the gradient magnitude)."""
    lib = _lib.load()
    _lib.require_cuda(images, "images")
    if images.dim() != 4 or images.shape[-1] != 3 or images.dtype not in (torch.float32, torch.uint8):
        raise ValueError(f"images must be [B,H,W,3] float32 or uint8, got {tuple(images.shape)} {images.dtype}")
    images = images.contiguous()
    B, H, W, _ = images.shape
    dev = images.device
    edges = out if out is not None else torch.empty((B, H, W), dtype=torch.uint8, device=dev)
    ws_bytes = lib.cm_canny_workspace_bytes(B, H, W)
    ws = WORKSPACE.get(dev, "canny", ws_bytes)
    _lib.check(lib.cm_canny(_lib.ptr(images), 1 if images.dtype == torch.uint8 else 0, B, H, W, float(sigma), float(low_threshold),
                            float(high_threshold), _lib.ptr(ws), ws_bytes, _lib.ptr(edges), _lib.stream_ptr(dev)), "cm_canny")
    return edges


def slic_segments(images_u8, n_segments: int = 500, compactness: float = 10.0, sigma: float = 1.0, max_num_iter: int = 10, out=None):
    """``skimage.segmentation.slic(img_u8, n_segments, compactness, sigma)`` of every image
    (extract_rg_embeddings.py:143-144) on the device: ``images_u8`` ``[B,H,W,3]`` uint8 -> ``[B,H,W]`` int32 labels
    (from 1, connectivity enforced).  ``cm_slic``; see oracle/slic.py for the restated algorithm (parity unpinned)."""
    lib = _lib.load()
    _lib.require_cuda(images_u8, "images")
    if images_u8.dim() != 4 or images_u8.shape[-1] != 3 or images_u8.dtype != torch.uint8:
        raise ValueError(f"images must be [B,H,W,3] uint8, got {tuple(images_u8.shape)} {images_u8.dtype}")
    images_u8 = images_u8.contiguous()
    B, H, W, _ = images_u8.shape
    dev = images_u8.device
    labels = out if out is not None else torch.empty((B, H, W), dtype=torch.int32, device=dev)
    ws_bytes = lib.cm_slic_workspace_bytes(B, H, W, int(n_segments))
    ws = WORKSPACE.get(dev, "slic", ws_bytes)
    _lib.check(lib.cm_slic(_lib.ptr(images_u8), B, H, W, int(n_segments), float(compactness), float(sigma), int(max_num_iter),
                           _lib.ptr(ws), ws_bytes, _lib.ptr(labels), _lib.stream_ptr(dev)), "cm_slic")
    return labels


class _CsrData(Data):
    """``Data`` that also carries the target-keyed CSR the build produced, so the GNN need not rebuild it."""


def create_region_graph(image, n_segments=500, segments=None, edges=None, device=None):
    """Drop-in for ``create_region_graph(image, n_segments) -> (Data, segments)`` (extract_rg_embeddings.py:138).

    ``image``: ``[H,W,3]`` float array in [0,1].  ``segments`` (the SLIC label map, :143-144) is computed on the device
    by ``slic_segments`` (``cm_slic``) from ``(image * 255).astype(uint8)`` like the reference unless given, ``edges``
    (the boolean Canny map, :152) by ``canny_edges`` (``cm_canny``, sigma 2) unless given.  The returned ``Data`` lives on the CUDA
    device; ``edge_index`` is in canonical (sorted) order and ``edge_attr`` is ``[E,1]``.
    """
    dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
    img = torch.as_tensor(np.asarray(image), dtype=torch.float32).to(dev)[None]
    if segments is None:
        img_u8 = torch.as_tensor((np.asarray(image) * 255).astype(np.uint8)).to(dev)[None]        # :143
        lab = slic_segments(img_u8, n_segments=n_segments, compactness=10.0, sigma=1.0)         # :144
        segments = lab[0].cpu().numpy().astype(np.int64)
    seg_np = np.asarray(segments)
    lab = torch.as_tensor(seg_np.astype(np.int32)).to(dev)[None]
    can = canny_edges(img, sigma=2.0) if edges is None else torch.as_tensor(np.asarray(edges).astype(np.uint8)).to(dev)[None]
    cap = max(int(seg_np.max()) + 1, 1) if seg_np.size else 1
    g = build_region_graphs(img, lab, can, label_cap=cap, local_ids=True)
    n, e = g.counts()
    d = _CsrData(x=g.x[:n], edge_index=g.edge_index[:, :e],
                 edge_attr=g.edge_w[:e].unsqueeze(1) if e > 0 else torch.zeros((0, 1), device=dev))
    d._csr = (g.csr_row_ptr[: n + 1], g.csr_col[:e])
    d.id_of_label = g.id_of_label[0]
    return d, segments


# ---------------------------------------------------------------------------- the reference's single-image front end
IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)


def denormalised_image(image_u8):
    """The float image ``extract_embeddings_from_image`` hands to ``create_region_graph``
    (extract_rg_embeddings.py:263-276) from the decoded ``[H,W,3]`` uint8 array: torchvision ``ToTensor`` (float32
    ``/255``) + ``Normalize`` (float32 ``sub(mean).div(std)``), then ``* std + mean`` in float64 and ``clip(0, 1)``.
    The round trip is not the identity: ``(image * 255).astype(uint8)`` (:143) lands one below the decoder's byte
    wherever the float64 result falls just short of it, so SLIC sees those bytes -- reproduced here operation for
    operation (host code: eleven flops per pixel on a 256 x 256 image)."""
    u8 = np.array(image_u8, dtype=np.uint8, order="C")      # a writable copy
    t = torch.from_numpy(u8).permute(2, 0, 1).contiguous().to(torch.float32).div(255)
    mean = torch.tensor(IMAGENET_MEAN, dtype=torch.float32).view(3, 1, 1)
    std = torch.tensor(IMAGENET_STD, dtype=torch.float32).view(3, 1, 1)
    a = t.sub(mean).div(std).numpy().transpose(1, 2, 0)
    return np.clip(a * np.array(IMAGENET_STD) + np.array(IMAGENET_MEAN), 0, 1)


def load_image_256(image_path):
    """``Image.open(path).convert('RGB').resize((256, 256))`` (:260-261) -> ``[256,256,3]`` uint8.  Decode and resample
    stay on the host (PIL), as in the reference."""
    from PIL import Image
    return np.asarray(Image.open(image_path).convert("RGB").resize((256, 256)), dtype=np.uint8)


def extract_embeddings_from_image(model, image_path, device, n_segments=500):
    """Drop-in for ``extract_embeddings_from_image(model, image_path, device, n_segments=500)``
    (extract_rg_embeddings.py:249-288): -> ``(node_embeddings [N,128] cpu, graph_embedding [1,128] cpu, graph_data cpu,
    segments [256,256] numpy)``.  The reference's ``predict_single_image`` (models/multimodal/test_multimodal.py:83-106)
    runs on top of it unchanged.  SLIC, Canny, the graph build and the GNN run on ``device`` (which must be a CUDA
    device: there is no CPU path); the GNN runs once for both embeddings (the reference runs it twice, :285-286)."""
    dev = torch.device(device)
    if dev.type != "cuda":
        raise _lib.CmError("extract_embeddings_from_image: the B200 path needs a CUDA device, got %r" % (device,))
    img_np = denormalised_image(load_image_256(image_path))
    graph_data, segments = create_region_graph(img_np, n_segments=n_segments, device=dev)
    graph_data = graph_data.to(dev)
    model.eval()                                                 # :283 (left in eval mode, as the reference does)
    with torch.no_grad():
        r = model._run_data(graph_data, ("node_emb", "graph_emb"))
    return r["node_emb"].cpu(), r["graph_emb"].cpu(), graph_data.cpu(), segments


def batch_to_data(g: RegionGraphBatch, counts: tuple[int, int] | None = None) -> Batch:
    """Exact-size PyG-style ``Batch`` view of a batched build (one device->host read for the sizes, unless the caller
    already holds them: ``counts`` = (nodes, edges), as ``RegionGraphPrefetcher`` does)."""
    n, e = counts if counts is not None else g.counts()
    out = Batch(x=g.x[:n], edge_index=g.edge_index[:, :e] if g.edge_index is not None else None,
                edge_attr=g.edge_w[:e].unsqueeze(1))
    out.batch = g.batch_vec[:n]
    out.ptr = g.node_ptr.to(torch.long)
    out.num_graphs = g.batch
    out._csr = (g.csr_row_ptr[: n + 1], g.csr_col[:e])
    return out


class RegionGraphPrefetcher:
    """The region graphs of the NEXT batch, built on a side stream while the current step's kernels run.

    A training step needs the node / edge counts on the host (exact-size views, kernel grids), and that read drains the
    GPU when the build sits at the top of the step: at batch 16 it cost 1 ms of a 6 ms step.  This is the overlap the
    reference gets from its ``DataLoader`` workers (models/region_graph/train.py builds graphs in ``__getitem__``):
    ``submit`` launches ``build_region_graphs`` + an asynchronous copy of (status, nodes, edges) into pinned memory on
    the prefetcher's stream; ``get`` waits for THAT copy only, orders the caller's stream behind the build and returns
    ``(RegionGraphBatch, Batch)`` -- the caller's stream never runs dry::

        pf = RegionGraphPrefetcher(label_cap=544)
        pf.submit(*first_batch)
        for batch in rest:
            g, data = pf.get()
            pf.submit(*batch)          # runs next to this step's kernels
            train_step(g, data)

    Inputs must stay unmodified until the matching ``get``.  ``ready``: an event after which the inputs are complete
    (default: everything queued on the caller's stream at ``submit``).  Build overflow raises in ``get``."""

    def __init__(self, device=None, **build_kwargs):
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        if build_kwargs.get("label_cap") is None:
            raise ValueError("RegionGraphPrefetcher needs label_cap (the default reads labels.max() back: a stall per batch)")
        if build_kwargs.get("out") is not None:
            raise ValueError("RegionGraphPrefetcher allocates a graph per batch (several are in flight)")
        self.build_kwargs = build_kwargs
        self.stream = torch.cuda.Stream(self.device)
        self._queue = []

    def __len__(self):
        return len(self._queue)

    def submit(self, images, labels, canny, ready: "torch.cuda.Event | None" = None):
        main = torch.cuda.current_stream(self.device)
        if ready is not None:
            self.stream.wait_event(ready)
        else:
            self.stream.wait_stream(main)
        with torch.cuda.stream(self.stream):
            g = build_region_graphs(images, labels, canny, **self.build_kwargs)
            host = torch.empty(3, dtype=torch.int32, pin_memory=True)
            host.copy_(torch.cat([g.status.view(1), g.node_ptr[-1:], g.edge_ptr[-1:]]), non_blocking=True)
            done = torch.cuda.Event()
            done.record(self.stream)
        self._queue.append((g, host, done, (images, labels, canny)))

    def get(self):
        if not self._queue:
            raise RuntimeError("RegionGraphPrefetcher.get() without a submitted batch")
        g, host, done, _inputs = self._queue.pop(0)
        done.synchronize()
        main = torch.cuda.current_stream(self.device)
        main.wait_event(done)
        for v in vars(g).values():      # allocated on the side stream, consumed on the caller's: reuse waits for the latter
            if isinstance(v, torch.Tensor):
                v.record_stream(main)
        s, n, e = (int(v) for v in host.tolist())
        if s:
            raise _lib.CmError(
                f"region-graph build overflow (status {s}): "
                + ", ".join(name for bit, name in ((1, "label >= label_cap"), (2, "nodes > node_cap"), (4, "edges > edge_cap")) if s & bit))
        return g, batch_to_data(g, (n, e))


# ------------------------------------------------------------------------------------------------ GNN
class _GATConvParams(nn.Module):
    """Parameter container with PyG >= 2.5 ``GATConv`` names (lin.weight, att_src, att_dst, bias)."""

    def __init__(self, in_channels, out_channels, heads):
        super().__init__()
        self.heads, self.out_channels = heads, out_channels
        self.lin = nn.Linear(in_channels, heads * out_channels, bias=False)
        self.att_src = nn.Parameter(torch.empty(1, heads, out_channels))
        self.att_dst = nn.Parameter(torch.empty(1, heads, out_channels))
        self.bias = nn.Parameter(torch.zeros(out_channels))
        for t in (self.lin.weight, self.att_src, self.att_dst):   # PyG glorot
            bound = (6.0 / (t.size(-2) + t.size(-1))) ** 0.5
            nn.init.uniform_(t, -bound, bound)


class _GCNConvParams(nn.Module):
    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.lin = nn.Linear(in_channels, out_channels, bias=False)
        self.bias = nn.Parameter(torch.zeros(out_channels))
        bound = (6.0 / (in_channels + out_channels)) ** 0.5
        nn.init.uniform_(self.lin.weight, -bound, bound)


class RegionGraphGNN(nn.Module):
    """``RegionGraphGNN(in_channels=15, hidden_channels=128, num_classes=2)`` (extract_rg_embeddings.py:27-135).

    Inference (eval mode) runs on the sm_100a kernels through ``cm_gnn_forward``; under ``model.train()``,
    ``forward`` runs ``cm_gnn_train_forward`` and is differentiable through ``cm_gnn_backward`` (training.py).
    There is no eager path.
    """

    def __init__(self, in_channels=15, hidden_channels=128, num_classes=2):
        super().__init__()
        h = hidden_channels
        self.in_channels, self.hidden_channels, self.num_classes = in_channels, h, num_classes
        self.conv1 = _GATConvParams(in_channels, h, heads=4)
        self.bn1 = nn.BatchNorm1d(h)
        self.conv2 = _GCNConvParams(h, h)
        self.bn2 = nn.BatchNorm1d(h)
        self.conv3 = _GCNConvParams(h, h)
        self.bn3 = nn.BatchNorm1d(h)
        self.conv4 = _GCNConvParams(h, h)
        self.bn4 = nn.BatchNorm1d(h)
        self.fc_shared = nn.Linear(h, h)
        self.fc_mask_1 = nn.Linear(h, h // 2)
        self.fc_mask_2 = nn.Linear(h // 2, num_classes)
        self.fc_instance_1 = nn.Linear(h, h // 2)
        self.fc_instance_2 = nn.Linear(h // 2, num_classes)
        self.fc_edge_1 = nn.Linear(h, h // 2)
        self.fc_edge_2 = nn.Linear(h // 2, 1)
        self._folded = None
        self._folded_key = None

    # checkpoints written with PyG 2.3/2.4 carry lin_src/lin_dst (the same tensor twice) instead of lin
    def load_state_dict(self, state_dict, strict=True, **kw):
        sd = dict(state_dict)
        if "conv1.lin.weight" not in sd and "conv1.lin_src.weight" in sd:
            sd["conv1.lin.weight"] = sd.pop("conv1.lin_src.weight")
            sd.pop("conv1.lin_dst.weight", None)
        return super().load_state_dict(sd, strict=strict, **kw)

    # ---- parameter marshalling
    def _params_struct(self):
        keep = []

        def P(t):
            t = _lib.f32c(t)
            keep.append(t)
            return t.data_ptr()

        s = _lib.GnnParams()
        s.in_channels, s.hidden, s.heads, s.num_classes = self.in_channels, self.hidden_channels, 4, self.num_classes
        s.conv1_lin_weight = P(self.conv1.lin.weight)
        s.conv1_att_src = P(self.conv1.att_src)
        s.conv1_att_dst = P(self.conv1.att_dst)
        s.conv1_bias = P(self.conv1.bias)
        for k, conv in enumerate((self.conv2, self.conv3, self.conv4)):
            s.conv_lin_weight[k] = P(conv.lin.weight)
            s.conv_bias[k] = P(conv.bias)
        for k, bn in enumerate((self.bn1, self.bn2, self.bn3, self.bn4)):
            s.bn_weight[k], s.bn_bias[k] = P(bn.weight), P(bn.bias)
            s.bn_mean[k], s.bn_var[k] = P(bn.running_mean), P(bn.running_var)
        s.bn_eps = float(self.bn1.eps)
        s.fc_shared_weight, s.fc_shared_bias = P(self.fc_shared.weight), P(self.fc_shared.bias)
        for k, (a, b) in enumerate(((self.fc_mask_1, self.fc_mask_2), (self.fc_instance_1, self.fc_instance_2),
                                    (self.fc_edge_1, self.fc_edge_2))):
            s.head1_weight[k], s.head1_bias[k] = P(a.weight), P(a.bias)
            s.head2_weight[k], s.head2_bias[k] = P(b.weight), P(b.bias)
        return s, keep

    def _prepared(self, device):
        """(params struct, folded blob); refolded whenever a parameter or buffer changed."""
        tensors = list(self.parameters()) + [b for b in self.buffers()]
        key = tuple((t.data_ptr(), t._version) for t in tensors) + (str(device), _lib.param_generation())
        if self._folded_key != key:
            lib = _lib.load()
            for t in tensors:
                _lib.require_cuda(t, "RegionGraphGNN parameter")
            s, keep = self._params_struct()
            nbytes = lib.cm_gnn_folded_bytes(C.byref(s))
            if nbytes == 0:
                _lib.check(lib.cm_gnn_fold(C.byref(s), None, 0, None), "cm_gnn_fold")
            blob = torch.empty(nbytes, dtype=torch.uint8, device=device)
            _lib.check(lib.cm_gnn_fold(C.byref(s), _lib.ptr(blob), nbytes, _lib.stream_ptr(device)), "cm_gnn_fold")
            self._folded, self._folded_key = (s, keep, blob), key
        return self._folded

    # ---- the one kernel-backed forward
    def run(self, x, csr_row_ptr, csr_col, edge_w=None, self_w=None, n_nodes_dev=None, node_ptr=None, n_graphs=0,
            want=("node_emb", "mask_out", "instance_out", "edge_out"), out=None):
        """Run A5-A7 on device buffers.  ``x`` may be capacity-sized with the live count in ``n_nodes_dev``."""
        if self.training:
            raise NotImplementedError("RegionGraphGNN on the B200 path is inference-only for now: call .eval()")
        lib = _lib.load()
        _lib.require_cuda(x, "data.x")
        dev = x.device
        s, _keep, blob = self._prepared(dev)
        cap = x.shape[0]
        h, c = self.hidden_channels, self.num_classes
        shapes = {"node_emb": (cap, h), "mask_out": (cap, c), "instance_out": (cap, c), "edge_out": (cap, 1),
                  "mask_prob": (cap,), "graph_emb": (max(n_graphs, 1), h)}
        res = {}
        for k in want:
            res[k] = out[k] if out is not None and k in out else torch.empty(shapes[k], dtype=torch.float32, device=dev)
        if cap == 0:
            if "graph_emb" in res:
                res["graph_emb"].zero_()
            return res
        ws_bytes = lib.cm_gnn_workspace_bytes(C.byref(s), cap)
        ws = WORKSPACE.get(dev, "gnn", ws_bytes)
        p = _lib.ptr
        x = _lib.f32c(x)
        _lib.check(lib.cm_gnn_forward(C.byref(s), p(blob), p(x), p(csr_row_ptr), p(csr_col), p(edge_w), p(self_w),
                                      p(n_nodes_dev), cap, p(node_ptr), int(n_graphs), p(ws), ws_bytes,
                                      p(res.get("node_emb")), p(res.get("mask_out")), p(res.get("instance_out")),
                                      p(res.get("edge_out")), p(res.get("mask_prob")), p(res.get("graph_emb")),
                                      _lib.stream_ptr(dev)), "cm_gnn_forward")
        return res

    def _csr_of(self, data):
        """Target-keyed CSR (+ weights) of a PyG-style ``data``; reuses the build's CSR when present."""
        x = data.x
        n = x.shape[0]
        ew = data.edge_attr.reshape(-1) if data.edge_attr is not None and data.edge_attr.numel() > 0 else None   # :56
        csr = getattr(data, "_csr", None)
        if csr is not None and csr[0].shape[0] == n + 1:
            return csr[0], csr[1], (_lib.f32c(ew) if ew is not None else None), None
        lib = _lib.load()
        ei = data.edge_index
        _lib.require_cuda(ei, "data.edge_index")
        ei = ei.to(torch.int64).contiguous()
        e = ei.shape[1]
        dev = x.device
        row_ptr = torch.empty(n + 1, dtype=torch.int32, device=dev)
        col = torch.empty(max(e, 1), dtype=torch.int32, device=dev)
        w = torch.empty(max(e, 1), dtype=torch.float32, device=dev) if ew is not None else None
        self_w = torch.empty(max(n, 1), dtype=torch.float32, device=dev)
        status = torch.zeros(1, dtype=torch.int32, device=dev)
        ws_bytes = lib.cm_csr_workspace_bytes(n, e)
        ws = WORKSPACE.get(dev, "csr", ws_bytes)
        p = _lib.ptr
        ewc = _lib.f32c(ew) if ew is not None else None
        _lib.check(lib.cm_csr_from_edge_index(p(ei), e, p(ewc), n, e, p(ws), ws_bytes, p(row_ptr), p(col), p(w),
                                              p(self_w), p(status), _lib.stream_ptr(dev)), "cm_csr_from_edge_index")
        if int(status.item()):
            raise IndexError("edge_index refers to a node id outside [0, num_nodes)")
        return row_ptr, col, w, self_w

    def _run_data(self, data, want):
        row_ptr, col, w, self_w = self._csr_of(data)
        node_ptr, n_graphs = None, 0
        if "graph_emb" in want:
            if hasattr(data, "batch") and data.batch is not None:   # :129
                n_graphs = int(getattr(data, "num_graphs", 0)) or (int(data.batch.max().item()) + 1 if data.x.shape[0] else 0)
                cnt = torch.bincount(data.batch, minlength=n_graphs)
                node_ptr = torch.zeros(n_graphs + 1, dtype=torch.int32, device=data.x.device)
                node_ptr[1:] = torch.cumsum(cnt, 0).to(torch.int32)
            else:
                n_graphs = 1
                node_ptr = torch.tensor([0, data.x.shape[0]], dtype=torch.int32, device=data.x.device)
        return self.run(data.x, row_ptr, col, w, self_w, None, node_ptr, n_graphs, want=want)

    # F.dropout probabilities of the reference's forward (extract_rg_embeddings.py:61-85); training mode only
    p_conv, p_fc = 0.3, 0.2

    def forward(self, data):
        if self.training:   # model.train(): batch-statistics BatchNorm, dropout, autograd through cm_gnn_backward
            from .training import gnn_train_forward
            return gnn_train_forward(self, data)[:3]
        r = self._run_data(data, ("mask_out", "instance_out", "edge_out"))
        return r["mask_out"], r["instance_out"], r["edge_out"]

    def extract_node_embeddings(self, data):
        return self._run_data(data, ("node_emb",))["node_emb"]

    def extract_graph_embedding(self, data):
        return self._run_data(data, ("node_emb", "graph_emb"))["graph_emb"]
