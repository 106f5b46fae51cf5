"""Training-mode passes behind the reference's ``model(batch)`` / ``loss.backward()`` step.

The reference trains with plain PyTorch autograd (models/region_graph/train.py:345-357,
models/multimodal/train_multimodal.py:232-275).  Here the forward under ``model.train()`` and its backward are the
library's ``cm_*_train_forward`` / ``cm_*_backward`` kernels, exposed to autograd as one ``torch.autograd.Function``
per model so that the reference's loop -- losses, ``clip_grad_norm_``, ``AdamW`` -- runs unchanged on top.
PyTorch supplies the parameter / gradient tensors, the dropout keep masks (its Philox stream cannot be reproduced
inside a kernel, so masks are data) and the BatchNorm running-statistics update; no forward or backward arithmetic
runs in PyTorch.
"""
from __future__ import annotations

import ctypes as C
import math
import os

import torch

from . import _lib


def _gnn_param_list(m):
    """Trainable tensors in the field order of ``cm_gnn_grads``."""
    convs = (m.conv2, m.conv3, m.conv4)
    bns = (m.bn1, m.bn2, m.bn3, m.bn4)
    h1 = (m.fc_mask_1, m.fc_instance_1, m.fc_edge_1)
    h2 = (m.fc_mask_2, m.fc_instance_2, m.fc_edge_2)
    return ([m.conv1.lin.weight, m.conv1.att_src, m.conv1.att_dst, m.conv1.bias]
            + [c.lin.weight for c in convs] + [c.bias for c in convs]
            + [b.weight for b in bns] + [b.bias for b in bns]
            + [m.fc_shared.weight, m.fc_shared.bias]
            + [l.weight for l in h1] + [l.bias for l in h1] + [l.weight for l in h2] + [l.bias for l in h2])


def _gnn_grads_struct(grads):
    g = _lib.GnnGrads()
    it = iter(t.data_ptr() for t in grads)
    g.conv1_lin_weight, g.conv1_att_src, g.conv1_att_dst, g.conv1_bias = next(it), next(it), next(it), next(it)
    for k in range(3):
        g.conv_lin_weight[k] = next(it)
    for k in range(3):
        g.conv_bias[k] = next(it)
    for k in range(4):
        g.bn_weight[k] = next(it)
    for k in range(4):
        g.bn_bias[k] = next(it)
    g.fc_shared_weight, g.fc_shared_bias = next(it), next(it)
    for arr in (g.head1_weight, g.head1_bias, g.head2_weight, g.head2_bias):
        for k in range(3):
            arr[k] = next(it)
    return g


def _dropout_struct(masks, p_conv, p_fc):
    d = _lib.GnnDropout()
    for k in range(7):
        d.keep[k] = masks[k].data_ptr() if masks[k] is not None else None
    d.p_conv, d.p_fc = float(p_conv), float(p_fc)
    return d


class GnnTrainFunction(torch.autograd.Function):
    """``(mask_out, instance_out, edge_out, node_emb) = f(model, x, csr..., masks, *params)``."""

    @staticmethod
    def forward(ctx, model, x, row_ptr, col, w, self_w, n_edges, masks, *params):
        lib = _lib.load()
        dev = x.device
        _lib.require_cuda(x, "data.x")
        x = _lib.f32c(x)
        n = x.shape[0]
        h, c = model.hidden_channels, model.num_classes
        held = [_lib.f32c(p) for p in params]
        s, keep = model._params_struct()
        drop = _dropout_struct(masks, model.p_conv, model.p_fc)
        ws_bytes = lib.cm_gnn_train_workspace_bytes(C.byref(s), n, n_edges)
        if ws_bytes == 0:
            raise ValueError("cm_gnn_train_workspace_bytes: " + lib.cm_last_error().decode())
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        f32 = dict(dtype=torch.float32, device=dev)
        mask_out, inst_out = torch.empty((n, c), **f32), torch.empty((n, c), **f32)
        edge_out, emb = torch.empty((n, 1), **f32), torch.empty((n, h), **f32)
        bmean, bvar = torch.empty((4, h), **f32), torch.empty((4, h), **f32)
        status = torch.zeros(1, dtype=torch.int32, device=dev)
        p = _lib.ptr
        _lib.check(lib.cm_gnn_train_forward(C.byref(s), p(x), p(row_ptr), p(col), p(w), p(self_w), n, n_edges, C.byref(drop),
                                            p(ws), ws_bytes, p(mask_out), p(inst_out), p(edge_out), p(emb), p(bmean), p(bvar),
                                            p(status), _lib.stream_ptr(dev)), "cm_gnn_train_forward")
        ctx.model, ctx.n_edges, ctx.masks = model, n_edges, masks
        ctx.params_struct = s      # the same tensors serve the backward (autograd forbids changing them in between)
        ctx.held = (x, row_ptr, col, w, self_w, ws, keep, held, status)
        sync = BatchNormSync.active
        if sync is not None:
            sync.raise_pending()
        # running statistics use the number of rows the statistics were taken over: every rank's under BatchNormSync
        update_bn_running_stats(model, bmean, bvar, sync.global_rows(h) if sync is not None else n)
        return mask_out, inst_out, edge_out, emb

    @staticmethod
    def backward(ctx, d_mask, d_inst, d_edge, d_emb):
        lib = _lib.load()
        model = ctx.model
        x, row_ptr, col, w, self_w, ws, keep, held, status = ctx.held
        dev = x.device
        # the symmetry flag of the forward's reverse-edge table: read without stalling the step (the backward is memory-safe
        # either way, its gradients are not) -- raised by the next training forward, or by check_deferred_status(block=True)
        _defer_status_check(status)
        s = ctx.params_struct
        drop = _dropout_struct(ctx.masks, model.p_conv, model.p_fc)
        grads = [torch.empty_like(t) for t in held]
        g = _gnn_grads_struct(grads)
        ds = [None if d is None else _lib.f32c(d) for d in (d_mask, d_inst, d_edge, d_emb)]
        p = _lib.ptr
        _lib.check(lib.cm_gnn_backward(C.byref(s), p(x), p(row_ptr), p(col), p(w), p(self_w), x.shape[0], ctx.n_edges,
                                       C.byref(drop), p(ws), ws.numel(), p(ds[0]), p(ds[1]), p(ds[2]), p(ds[3]), C.byref(g),
                                       _lib.stream_ptr(dev)), "cm_gnn_backward")
        if BatchNormSync.active is not None:
            BatchNormSync.active.raise_pending()
        return (None,) * 8 + tuple(grads)


def keep_masks(shapes, probs, device):
    """Dropout keep masks (uint8, 1 = keep with probability 1 - p) for a list of sites: ``None`` where the shape is
    ``None`` or p <= 0.  All sites with the same p are drawn by ONE ``torch.rand`` + ONE compare (the masks are views of
    one byte buffer, each starting on a 16-byte boundary) -- two launches per distinct p instead of three per site."""
    out = [None] * len(shapes)
    by_p = {}
    for i, (sh, p) in enumerate(zip(shapes, probs)):
        if sh is not None and p > 0:
            by_p.setdefault(float(p), []).append(i)
    for p, idx in by_p.items():
        offs, total = [], 0
        for i in idx:
            offs.append(total)
            total += (math.prod(shapes[i]) + 15) // 16 * 16
        keep = (torch.rand(total, device=device) >= p).view(torch.uint8)
        for i, o in zip(idx, offs):
            out[i] = keep[o:o + math.prod(shapes[i])].view(shapes[i])
    return out


_pending_status = []   # (event, pinned int32[1]) of backward passes whose symmetry flag has not been looked at yet
_ASYMMETRIC = "RegionGraphGNN training needs a structurally symmetric edge_index (every region adjacency is)"


def _defer_status_check(status):
    if os.environ.get("CM_TRAIN_EAGER_STATUS"):     # the stalling form: the error surfaces inside this backward
        if int(status.item()) & _lib.STATUS_EDGE_OVERFLOW:
            raise _lib.CmError(_ASYMMETRIC)
        return
    host = torch.empty(1, dtype=torch.int32, pin_memory=True)
    host.copy_(status, non_blocking=True)
    ev = torch.cuda.Event()
    ev.record(torch.cuda.current_stream(status.device))
    _pending_status.append((ev, host))


def check_deferred_status(block: bool = False):
    """Raises if a finished GNN backward ran on an edge_index that was not structurally symmetric (its gradients are
    wrong).  ``block=True`` waits for every backward still in flight -- call it before trusting the last step of a loop;
    every training forward calls the non-blocking form."""
    bad, keep = False, []
    for ev, host in _pending_status:
        if block:
            ev.synchronize()
        if ev.query():
            bad = bad or bool(int(host[0]) & _lib.STATUS_EDGE_OVERFLOW)
        else:
            keep.append((ev, host))
    _pending_status[:] = keep
    if bad:
        raise _lib.CmError(_ASYMMETRIC)


def gnn_train_forward(model, data):
    """``RegionGraphGNN.forward`` under ``model.train()``: returns ``(mask_out, instance_out, edge_out, node_emb)``
    attached to autograd and updates the BatchNorm running statistics (momentum 0.1, unbiased variance) like ``nn.BatchNorm1d`` does."""
    check_deferred_status()
    row_ptr, col, w, self_w = model._csr_of(data)
    x = data.x
    n = x.shape[0]
    if n == 0:
        raise ValueError("RegionGraphGNN training step on an empty graph")
    if getattr(data, "_csr", None) is not None and data._csr[0] is row_ptr:
        n_edges = int(col.shape[0])            # batch_to_data cut the build's CSR to its edge count: no second read
    else:
        n_edges = int(row_ptr[-1].item()) if row_ptr.shape[0] == n + 1 else 0
    h = model.hidden_channels
    dev = x.device
    masks = keep_masks([(n, h if k < 4 else h // 2) for k in range(7)], [model.p_conv if k < 3 else model.p_fc for k in range(7)], dev)
    params = _gnn_param_list(model)
    return GnnTrainFunction.apply(model, x, row_ptr, col, w, self_w, n_edges, masks, *params)


def update_bn_running_stats(model, batch_mean, batch_var, n):
    """``nn.BatchNorm1d``'s running-statistics update (momentum 0.1, unbiased variance).  ``n``: rows the batch statistics
    were taken over -- a Python int, or a 0-dim device tensor (the all-reduced row count under ``BatchNormSync``)."""
    with torch.no_grad():
        if torch.is_tensor(n):
            n = n.to(torch.float32)
            unbias = n / torch.clamp(n - 1, min=1)
        else:
            unbias = n / max(n - 1, 1)
        bns = (model.bn1, model.bn2, model.bn3, model.bn4)
        moms = [bn.momentum if bn.momentum is not None else 0.1 for bn in bns]
        if len(set(moms)) == 1:     # the usual case: the same arithmetic as the loop below in 6 launches instead of 28
            m = moms[0]
            means, vars_ = [bn.running_mean for bn in bns], [bn.running_var for bn in bns]
            torch._foreach_mul_(means + vars_, 1 - m)
            torch._foreach_add_(means, list(batch_mean.unbind(0)), alpha=m)
            torch._foreach_add_(vars_, list((batch_var * unbias * m).unbind(0)))
            torch._foreach_add_([bn.num_batches_tracked for bn in bns], 1)
            return
        for k, bn in enumerate(bns):
            m = moms[k]
            bn.running_mean.mul_(1 - m).add_(batch_mean[k], alpha=m)
            bn.running_var.mul_(1 - m).add_(batch_var[k] * unbias * m)
            bn.num_batches_tracked += 1


def allreduce_bn_sums(sums: torch.Tensor, group=None):
    """SUM-all-reduce of a BatchNorm statistics vector -- ``(sum x [H], sum x^2 [H], n)`` in the forward, ``(sum dy [H],
    sum dy*xhat [H])`` in the backward, float64 -- over the ranks, in place.  The per-layer exchange that makes a sharded
    batch normalise like the single-process reference (``Batch.from_data_list`` + BatchNorm over all nodes, train.py:231,
    241-250).  No-op without an initialised process group."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=group)
    return sums


class BatchNormSync:
    """Installs the library's data-parallel BatchNorm hook (``cm_set_bn_sync``) for the current process: while active,
    every BatchNorm layer of ``cm_gnn_train_forward`` / ``cm_gnn_backward`` hands its float64 column sums to ``reduce``
    (default ``allreduce_bn_sums``: one small NCCL all-reduce per layer and direction, on the launching stream) and
    normalises with the reduced sums, so N ranks with B/N images each compute exactly what one process computes on B
    images.  Use as a context manager around the training step, or ``install()`` / ``remove()``.  One per process."""
    active = None

    def __init__(self, device, hidden: int = 128, reduce=None, group=None):
        self.buf = torch.zeros(2 * hidden + 1, dtype=torch.float64, device=device)
        self.reduce = reduce if reduce is not None else (lambda t: allreduce_bn_sums(t, group))
        self._pending = None
        self._cb = _lib.BN_SYNC_FN(self._call)          # keeps the trampoline alive as long as the object lives
        self.calls = 0

    def _call(self, _user, count, _stream):
        try:   # the library has queued the copy into self.buf on torch's current stream of this device
            self.calls += 1
            self.reduce(self.buf[:count])
        except BaseException as e:   # noqa: BLE001 -- a ctypes callback cannot propagate: re-raised after the library call
            self._pending = e

    def raise_pending(self):
        e, self._pending = self._pending, None
        if e is not None:
            raise e

    def global_rows(self, hidden):
        """0-dim device tensor: the all-reduced row count of the last forward."""
        return self.buf[2 * hidden].clone()

    def install(self):
        _lib.check(_lib.load().cm_set_bn_sync(C.cast(self._cb, C.c_void_p), None, _lib.ptr(self.buf), self.buf.numel()), "cm_set_bn_sync")
        BatchNormSync.active = self
        return self

    def remove(self):
        _lib.check(_lib.load().cm_set_bn_sync(None, None, None, 0), "cm_set_bn_sync")
        if BatchNormSync.active is self:
            BatchNormSync.active = None

    __enter__ = install

    def __exit__(self, *exc):
        self.remove()
        return False


# ------------------------------------------------------------------------------------------------ fusion
def _fusion_param_list(runner):
    """Trainable tensors behind a ``_FusionRunner`` (a detector, or a bare CrossAttentionFusion / LateFusion) in
    ``cm_fusion_grads`` field order: the 28 cross-attention tensors, the 16 head tensors, the 6 late-fusion tensors;
    ``None`` where the module has no such tensor (``nn.Identity`` projection, no heads, the other fusion type)."""
    import torch.nn as nn
    from .fusion_model import CrossAttentionFusion
    f = runner.fusion
    lin = lambda m: (m.weight, m.bias) if isinstance(m, nn.Linear) else (None, None)
    if isinstance(f, CrossAttentionFusion):
        out = [*lin(f.rg_proj), *lin(f.kg_proj)]
        for m in (f.cross_attn_rg2kg, f.cross_attn_kg2rg):
            out += [m.in_proj_weight, m.in_proj_bias, m.out_proj.weight, m.out_proj.bias]
        out += [f.ln_rg.weight, f.ln_rg.bias, f.ln_kg.weight, f.ln_kg.bias]
        for seq in (f.ffn_rg, f.ffn_kg, f.fusion_layer):
            out += [seq[0].weight, seq[0].bias, seq[3].weight, seq[3].bias]
        late = [None] * 6
    else:
        out = [None] * len(_lib.FUSION_GRAD_FIELDS)
        late = [f.fusion[i].weight for i in (0, 3, 6)] + [f.fusion[i].bias for i in (0, 3, 6)]
    heads = runner.heads
    if heads is not None:
        out += [h[0].weight for h in heads] + [h[0].bias for h in heads] + [h[3].weight for h in heads] + [h[3].bias for h in heads]
    else:
        out += [None] * 16
    return out + late


def _fusion_grads_struct(grads):
    g = _lib.FusionGrads()
    it = iter(grads)
    dp = lambda t: t.data_ptr() if t is not None else None
    for name in _lib.FUSION_GRAD_FIELDS:
        setattr(g, name, dp(next(it)))
    for arr in (g.head_w1, g.head_b1, g.head_w2, g.head_b2):
        for k in range(4):
            arr[k] = dp(next(it))
    for arr in (g.late_w, g.late_b):
        for k in range(3):
            arr[k] = dp(next(it))
    return g


def _fusion_dropout_struct(masks, p):
    d = _lib.FusionDropout()
    for k in range(9):
        d.keep[k] = masks[k].data_ptr() if masks is not None and masks[k] is not None else None
    d.p = float(p)
    return d


class FusionTrainFunction(torch.autograd.Function):
    """``(fused[, mask_out, instance_out, edge_out, score_out]) = f(runner, p_drop, rg_rows, node_ptr, G, kg, masks,
    *params)``; differentiable in ``rg_rows``, ``kg`` and the parameters.  ``runner``: the ``_FusionRunner`` of a
    detector (five outputs) or of a bare fusion module (``fused`` only); either fusion type."""

    @staticmethod
    def forward(ctx, runner, p_drop, rg_rows, node_ptr, n_graphs, kg, masks, *params):
        lib = _lib.load()
        _lib.require_cuda(rg_rows, "rg_embeddings")
        _lib.require_cuda(kg, "kg_embeddings")
        dev = rg_rows.device
        rg_rows, kg = _lib.f32c(rg_rows), _lib.f32c(kg)
        G, M, K = int(n_graphs), rg_rows.shape[0], kg.shape[1]
        if kg.shape[0] != G:
            raise RuntimeError(f"training step needs one KG copy per image: kg has {kg.shape[0]} graphs, batch has {G}")
        s, keep = runner.params()
        drop = _fusion_dropout_struct(masks, p_drop)
        ws_bytes = lib.cm_fusion_train_workspace_bytes(C.byref(s), G, M, K)
        if ws_bytes == 0:
            raise ValueError("cm_fusion_train_workspace_bytes: " + lib.cm_last_error().decode())
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        f32 = dict(dtype=torch.float32, device=dev)
        c = s.num_classes
        final = s.hidden if s.fusion_type == 0 else s.hidden // 2
        fused = torch.empty((G, final), **f32)
        has_heads = runner.heads is not None
        outs = ([torch.empty((G, c), **f32), torch.empty((G, c), **f32), torch.empty((G, 1), **f32), torch.empty((G, 1), **f32)]
                if has_heads else [None] * 4)
        p = _lib.ptr
        _lib.check(lib.cm_fusion_train_forward(C.byref(s), p(rg_rows), p(node_ptr), G, M, p(kg), K, C.byref(drop), p(ws), ws_bytes,
                                               p(fused), p(outs[0]), p(outs[1]), p(outs[2]), p(outs[3]), _lib.stream_ptr(dev)),
                   "cm_fusion_train_forward")
        ctx.runner, ctx.p_drop, ctx.masks, ctx.dims, ctx.has_heads = runner, p_drop, masks, (G, M, K), has_heads
        ctx.held = (rg_rows, node_ptr, kg, ws, keep, [None if t is None else _lib.f32c(t) for t in params])
        return (fused, *outs) if has_heads else fused

    @staticmethod
    def backward(ctx, d_fused, *d_heads):
        lib = _lib.load()
        rg_rows, node_ptr, kg, ws, keep, held = ctx.held
        G, M, K = ctx.dims
        dev = rg_rows.device
        s, keep2 = ctx.runner.params()
        drop = _fusion_dropout_struct(ctx.masks, ctx.p_drop)
        grads = [None if t is None else torch.empty_like(t) for t in held]
        g = _fusion_grads_struct(grads)
        d_heads = tuple(d_heads) + (None,) * (4 - len(d_heads))
        ds = [None if d is None else _lib.f32c(d) for d in (d_fused, *d_heads)]
        d_rg = torch.empty_like(rg_rows) if ctx.needs_input_grad[2] else None
        d_kg = torch.empty_like(kg) if ctx.needs_input_grad[5] else None
        p = _lib.ptr
        _lib.check(lib.cm_fusion_backward(C.byref(s), p(rg_rows), p(node_ptr), G, M, p(kg), K, C.byref(drop), p(ws), ws.numel(),
                                          p(ds[0]), p(ds[1]), p(ds[2]), p(ds[3]), p(ds[4]), C.byref(g), p(d_rg), p(d_kg),
                                          _lib.stream_ptr(dev)), "cm_fusion_backward")
        return (None, None, d_rg, None, None, d_kg, None) + tuple(grads)


def fusion_dropout_masks(runner, p_drop, G, M, K, device):
    """Fresh keep masks (uint8) for every dropout site of the module behind ``runner`` (``None`` when p = 0): cross
    attention -- attention rg->kg [M,H,K], FFN_rg [M,2D], attention kg->rg [H,K,M], FFN_kg [G*K,2D], fusion MLP [G,D],
    four head hidden layers [G,D/2]; late -- [G,D], [G,D/2], then the four head hidden layers [G,D/4] at 5..8."""
    if p_drop <= 0:
        return None
    from .fusion_model import CrossAttentionFusion
    f = runner.fusion
    D = f.hidden_dim
    if isinstance(f, CrossAttentionFusion):
        H = f.cross_attn_rg2kg.num_heads
        shapes = [(M, H, K), (M, 2 * D), (H, K, M), (G * K, 2 * D), (G, D)] + [(G, D // 2)] * 4
    else:
        shapes = [(G, D), (G, D // 2), None, None, None] + [(G, D // 4)] * 4
    if runner.heads is None:
        shapes = shapes[:5] + [None] * 4
    return keep_masks(shapes, [p_drop] * len(shapes), device)


def fusion_train_forward(det, rg_rows, node_ptr, n_graphs, kg, masks="auto"):
    """Training-mode forward on a ragged batch of a detector -> ``(fused, mask, instance, edge, score)``, or of a bare
    fusion module (``CrossAttentionFusion`` / ``LateFusion``) -> ``fused``."""
    runner = det._runner
    G, M, K = int(n_graphs), rg_rows.shape[0], kg.shape[1]
    if kg.shape[0] == 1 and G > 1:
        kg = kg.expand(G, -1, -1)
    kg = kg.contiguous()
    p_drop = det.dropout_p
    if masks == "auto":
        masks = fusion_dropout_masks(runner, p_drop, G, M, K, rg_rows.device)
    params = _fusion_param_list(runner)
    return FusionTrainFunction.apply(runner, p_drop, rg_rows, node_ptr, G, kg, masks, *params)


# ------------------------------------------------------------------------------------------------ data parallel
def allreduce_gradients(parameters, average: bool = False):
    """Sum (the reference's fusion step sums per-sample gradients, train_multimodal.py:270) or average the ``.grad``
    of every parameter over ranks in ONE flat all-reduce (6.2 MB of fp32 for both models: a single NCCL call over
    NVLink sized for launch latency, not one call per tensor).  No-op without an initialised process group."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return
    grads = [p.grad for p in parameters if p.grad is not None]
    if not grads:
        return
    flat = torch.cat([g.reshape(-1) for g in grads])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM)
    if average:
        flat /= dist.get_world_size()
    off = 0
    for g in grads:
        n = g.numel()
        g.copy_(flat[off:off + n].view_as(g))
        off += n


def joint_train_step(gnn, det, data, node_ptr, n_graphs, kg, loss_fn, optimizer=None, max_norm: float = 1.0,
                     average_gradients: bool = False):
    """One data-parallel training step through both models (BASELINE configs[3]): GNN forward (train mode) -> node
    embeddings -> fusion + heads -> ``loss_fn(node_outputs, image_outputs)`` -> backward through
    ``cm_fusion_backward`` and ``cm_gnn_backward`` -> gradient all-reduce -> ``clip_grad_norm_`` -> optimizer step
    (models/region_graph/train.py:345-357, models/multimodal/train_multimodal.py:232-275).  Returns the loss."""
    params = list(gnn.parameters()) + list(det.parameters())
    if optimizer is not None:
        optimizer.zero_grad(set_to_none=True)
    mask, inst, edge, emb = gnn_train_forward(gnn, data)
    res = det.forward_ragged(emb, node_ptr, n_graphs, kg)
    loss = loss_fn((mask, inst, edge), res)
    loss.backward()
    if hasattr(optimizer, "allreduce"):     # train_loop.FusedAdamW: gradients already live in one flat buffer
        optimizer.allreduce(average=average_gradients)
    else:
        allreduce_gradients(params, average=average_gradients)
    if max_norm is not None:
        torch.nn.utils.clip_grad_norm_(params, max_norm=max_norm)
    if optimizer is not None:
        optimizer.step()
    return loss
