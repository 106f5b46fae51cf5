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
        ctx.held = (x, row_ptr, col, w, self_w, ws, keep, held, status)
        update_bn_running_stats(model, bmean, bvar, n)
        return mask_out, inst_out, edge_out, emb

    @staticmethod
    def backward(ctx, d_mask, d_inst, d_edge, d_emb):
        lib = _lib.load()
        model = ctx.model
        x, row_ptr, col, w, self_w, ws, keep, held, status = ctx.held
        dev = x.device
        if int(status.item()) & _lib.STATUS_EDGE_OVERFLOW:
            raise _lib.CmError("RegionGraphGNN training needs a structurally symmetric edge_index (every region adjacency is)")
        s, keep2 = model._params_struct()
        drop = _dropout_struct(ctx.masks, model.p_conv, model.p_fc)
        grads = [torch.empty_like(t) for t in held]
        g = _gnn_grads_struct(grads)
        ds = [None if d is None else _lib.f32c(d) for d in (d_mask, d_inst, d_edge, d_emb)]
        p = _lib.ptr
        _lib.check(lib.cm_gnn_backward(C.byref(s), p(x), p(row_ptr), p(col), p(w), p(self_w), x.shape[0], ctx.n_edges,
                                       C.byref(drop), p(ws), ws.numel(), p(ds[0]), p(ds[1]), p(ds[2]), p(ds[3]), C.byref(g),
                                       _lib.stream_ptr(dev)), "cm_gnn_backward")
        return (None,) * 8 + tuple(grads)


def gnn_train_forward(model, data):
    """``RegionGraphGNN.forward`` under ``model.train()``: returns ``(mask_out, instance_out, edge_out, node_emb)``
    attached to autograd and updates the BatchNorm running statistics (momentum 0.1, unbiased variance) like ``nn.BatchNorm1d`` does."""
    row_ptr, col, w, self_w = model._csr_of(data)
    x = data.x
    n = x.shape[0]
    if n == 0:
        raise ValueError("RegionGraphGNN training step on an empty graph")
    n_edges = int(row_ptr[-1].item()) if row_ptr.shape[0] == n + 1 else 0
    h = model.hidden_channels
    dev = x.device
    masks = []
    for k in range(7):
        p_drop = model.p_conv if k < 3 else model.p_fc
        width = h if k < 4 else h // 2
        masks.append((torch.rand((n, width), device=dev) >= p_drop).to(torch.uint8) if p_drop > 0 else None)
    params = _gnn_param_list(model)
    return GnnTrainFunction.apply(model, x, row_ptr, col, w, self_w, n_edges, masks, *params)


def update_bn_running_stats(model, batch_mean, batch_var, n):
    with torch.no_grad():
        for k, bn in enumerate((model.bn1, model.bn2, model.bn3, model.bn4)):
            m = bn.momentum if bn.momentum is not None else 0.1
            bn.running_mean.mul_(1 - m).add_(batch_mean[k], alpha=m)
            bn.running_var.mul_(1 - m).add_(batch_var[k] * (n / max(n - 1, 1)), alpha=m)
            bn.num_batches_tracked += 1
