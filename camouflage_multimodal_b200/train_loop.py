"""The pieces of the reference's training loops that sit around the model passes, on the device:

* ``region_targets``       per-region labels from the ground-truth masks   (models/region_graph/train.py:169-176)
* ``RegionGraphLoss``      weighted CE + CE + BCE(pos_weight), task weights   (train.py:325-353)
* ``AggressiveFocalLoss`` / ``MultimodalLoss``   focal * 3 + CE + 0.5 BCE + 0.3 MSE   (train_multimodal.py:29-57, 253-268)
* ``FusedAdamW``           ``clip_grad_norm_`` + ``torch.optim.AdamW.step`` in two launches   (train.py:318, 356-357)

Names, defaults and reductions follow the reference; every value and gradient is computed by ``libcm_b200.so``
(``cm_region_targets``, ``cm_rg_node_loss``, ``cm_image_loss``, ``cm_adamw_step``) -- autograd only carries the
gradient tensors between the kernels.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib


def _stream(t=None):
    """Current stream of the device ``t`` lives on (the library call is then made with that device current)."""
    return _lib.stream_ptr(t.device if t is not None else None)


def region_targets(graph, labels, mask, instance_mask, edge_mask):
    """``graph``: RegionGraphBatch of the same ``labels`` [B,H,W] int32; masks [B,H,W] uint8 (the bytes PIL's
    ``convert('L')`` yields; the reference divides by 255 and thresholds the region mean at 0.5 / 0.5 / 0.3).
    Returns ``(y, instance_y, edge_y)``: int64, int64, float32, each ``[node_cap]`` (rows past the node count are 0)."""
    b, h, w = labels.shape
    for t in (mask, instance_mask, edge_mask):
        if t.dtype != torch.uint8 or t.shape != labels.shape or not t.is_cuda:
            raise ValueError("masks must be CUDA uint8 tensors shaped like labels")
    y = torch.zeros(graph.node_cap, dtype=torch.int64, device=labels.device)
    iy = torch.zeros_like(y)
    ey = torch.zeros(graph.node_cap, dtype=torch.float32, device=labels.device)
    lib = _lib.load()
    _lib.check(lib.cm_region_targets(_lib.ptr(labels.contiguous()), _lib.ptr(mask.contiguous()), _lib.ptr(instance_mask.contiguous()),
                                     _lib.ptr(edge_mask.contiguous()), b, h, w, _lib.ptr(graph.node_ptr), _lib.ptr(graph.node_label),
                                     _lib.ptr(graph.node_bbox), graph.node_cap, _lib.ptr(y), _lib.ptr(iy), _lib.ptr(ey), _stream(y)),
               "cm_region_targets")
    return y, iy, ey


class _RgLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mask, inst, edge, y, iy, ey, n_dev, cfg):
        cap = mask.shape[0]
        mask, inst, edge = mask.contiguous(), inst.contiguous(), edge.contiguous().view(-1)
        out = torch.empty(4, dtype=torch.float32, device=mask.device)
        ws = torch.empty(5, dtype=torch.float64, device=mask.device)
        dm, di, de = torch.empty_like(mask), torch.empty_like(inst), torch.empty_like(edge)
        _lib.check(_lib.load().cm_rg_node_loss(_lib.ptr(mask), _lib.ptr(inst), _lib.ptr(edge), _lib.ptr(y), _lib.ptr(iy), _lib.ptr(ey),
                                               _lib.ptr(n_dev) if n_dev is not None else None, cap, C.byref(cfg), _lib.ptr(ws), 40,
                                               _lib.ptr(out), _lib.ptr(dm), _lib.ptr(di), _lib.ptr(de), _stream(mask)), "cm_rg_node_loss")
        ctx.save_for_backward(dm, di, de)
        ctx.edge_shape = None
        ctx.terms = out
        return out[0]

    @staticmethod
    def backward(ctx, g):
        dm, di, de = ctx.saved_tensors
        return g * dm, g * di, g * de, None, None, None, None, None


class RegionGraphLoss(torch.nn.Module):
    """The multi-task criterion of models/region_graph/train.py:325-353 as one module (value + gradients from
    ``cm_rg_node_loss``).  ``terms`` holds (total, mask, instance, edge) of the last call on the device."""

    def __init__(self, mask_weight=(1.0, 5.0), instance_weight=(1.0, 4.0), edge_pos_weight=3.0, task_weights=None):
        super().__init__()
        tw = task_weights or {"mask": 2.0, "instance": 1.0, "edge": 0.5}
        if mask_weight[0] != 1.0 or instance_weight[0] != 1.0:
            raise ValueError("class-0 weights other than 1 are not supported (the reference uses [1, 5] and [1, 4])")
        self.cfg = _lib.RgLossConfig(mask_weight[1], instance_weight[1], edge_pos_weight, tw["mask"], tw["instance"], tw["edge"])
        self.terms = None

    def forward(self, mask_out, instance_out, edge_out, y, instance_y, edge_y, n_nodes_dev=None):
        edge_flat = edge_out.reshape(-1)
        loss = _RgLossFn.apply(mask_out, instance_out, edge_flat, y, instance_y, edge_y.to(torch.float32), n_nodes_dev, self.cfg)
        return loss


class _ImageLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mask, inst, edge, score, mask_label, edge_label, score_label, cfg):
        g = mask.shape[0]
        mask, inst = mask.contiguous(), inst.contiguous()
        edge, score = edge.contiguous().view(-1), score.contiguous().view(-1)
        out = torch.empty(1, dtype=torch.float32, device=mask.device)
        dm, di, de, ds = torch.empty_like(mask), torch.empty_like(inst), torch.empty_like(edge), torch.empty_like(score)
        _lib.check(_lib.load().cm_image_loss(_lib.ptr(mask), _lib.ptr(inst), _lib.ptr(edge), _lib.ptr(score), _lib.ptr(mask_label),
                                             _lib.ptr(edge_label), _lib.ptr(score_label), g, C.byref(cfg), _lib.ptr(out),
                                             _lib.ptr(dm), _lib.ptr(di), _lib.ptr(de), _lib.ptr(ds), _stream(mask)), "cm_image_loss")
        ctx.save_for_backward(dm, di, de, ds)
        return out[0]

    @staticmethod
    def backward(ctx, g):
        dm, di, de, ds = ctx.saved_tensors
        return g * dm, g * di, g * de, g * ds, None, None, None, None


class MultimodalLoss(torch.nn.Module):
    """``3 * AggressiveFocalLoss(0.75, 3)(mask) + CE(instance) + 0.5 * BCEWithLogits(edge) + 0.3 * MSE(score)`` per
    image, summed over the batch -- the reference calls ``loss.backward()`` once per sample, so the batch gradient is
    the SUM over images (models/multimodal/train_multimodal.py:253-268)."""

    def __init__(self, alpha=0.75, gamma=3.0, weights=(3.0, 1.0, 0.5, 0.3)):
        super().__init__()
        self.cfg = _lib.ImageLossConfig(alpha, gamma, *weights)

    def forward(self, mask_out, instance_out, edge_out, score_out, mask_label, edge_label, score_label):
        return _ImageLossFn.apply(mask_out, instance_out, edge_out.reshape(-1), score_out.reshape(-1), mask_label.to(torch.int64),
                                  edge_label.to(torch.float32), score_label.to(torch.float32), self.cfg)


class AggressiveFocalLoss(torch.nn.Module):
    """models/multimodal/train_multimodal.py:29-57: ``mean(alpha_t * (1 - p_t)^gamma * CE)`` (the other three terms of
    ``cm_image_loss`` get weight 0)."""

    def __init__(self, alpha=0.75, gamma=3.0):
        super().__init__()
        self.alpha, self.gamma = alpha, gamma

    def forward(self, inputs, targets):
        g = inputs.shape[0]
        cfg = _lib.ImageLossConfig(self.alpha, self.gamma, 1.0 / g, 0.0, 0.0, 0.0)
        z = torch.zeros(g, dtype=torch.float32, device=inputs.device)
        return _ImageLossFn.apply(inputs, inputs.detach(), z, z, targets.to(torch.int64), z, z, cfg)


class FusedAdamW:
    """``torch.nn.utils.clip_grad_norm_(params, max_norm)`` + ``torch.optim.AdamW(params, lr, weight_decay=1e-4).step()``
    (train.py:318, 356-357) as ``cm_adamw_step``: one launch for the squared norms of every tensor, one tiny launch for
    the clip coefficient, one launch for the update -- no host synchronisation, no per-tensor launches.
    ``total_norm`` (device scalar) is the gradient norm before clipping, as ``clip_grad_norm_`` returns it."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, max_norm=1.0):
        self.params = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError("no trainable parameters")
        dev = self.params[0].device
        for p in self.params:
            if p.dtype != torch.float32 or not p.is_cuda or not p.is_contiguous():
                raise ValueError("FusedAdamW takes contiguous float32 CUDA parameters")
        self.lr, self.betas, self.eps, self.weight_decay, self.max_norm = lr, betas, eps, weight_decay, max_norm
        self.exp_avg = [torch.zeros_like(p) for p in self.params]
        self.exp_avg_sq = [torch.zeros_like(p) for p in self.params]
        # stable gradient buffers, views of ONE flat tensor (zero_grad is one memset, a data-parallel all-reduce one call);
        # p.grad is pointed at them and autograd accumulates in place
        offs, total = [], 0
        for p in self.params:
            offs.append(total)
            total += (p.numel() + 3) // 4 * 4
        self.flat_grad = torch.zeros(total, dtype=torch.float32, device=dev)
        self.grads = [self.flat_grad[o:o + p.numel()].view_as(p) for o, p in zip(offs, self.params)]
        for p, g in zip(self.params, self.grads):
            p.grad = g
        table = lambda ts: torch.tensor([t.data_ptr() for t in ts], dtype=torch.int64, device=dev)  # noqa: E731
        self._p, self._g, self._m, self._v = table(self.params), table(self.grads), table(self.exp_avg), table(self.exp_avg_sq)
        self._numel = torch.tensor([p.numel() for p in self.params], dtype=torch.int64, device=dev)
        self._max_numel = max(p.numel() for p in self.params)
        self._norm_ws = torch.zeros(len(self.params), dtype=torch.float64, device=dev)
        self._norm_out = torch.zeros(2, dtype=torch.float32, device=dev)
        self.step_count = 0

    def zero_grad(self, set_to_none: bool = False):
        """Zeroes the registered buffers (``set_to_none`` is accepted for torch.optim compatibility and ignored: the
        kernels need stable gradient addresses)."""
        self.flat_grad.zero_()
        for p, g in zip(self.params, self.grads):
            p.grad = g

    def allreduce(self, average: bool = False):
        """Data-parallel gradient sum (or mean) over the process group: one NCCL call on the flat buffer."""
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            dist.all_reduce(self.flat_grad)
            if average:
                self.flat_grad.div_(dist.get_world_size())

    @property
    def total_norm(self):
        return self._norm_out[0]

    def step(self):
        for p, g in zip(self.params, self.grads):
            if p.grad is None:
                g.zero_()
                p.grad = g
            elif p.grad.data_ptr() != g.data_ptr():   # autograd replaced the tensor: copy into the registered buffer
                g.copy_(p.grad)
                p.grad = g
        self.step_count += 1
        cfg = _lib.AdamWConfig(self.lr, self.betas[0], self.betas[1], self.eps, self.weight_decay,
                               self.max_norm if self.max_norm is not None else 0.0)
        _lib.check(_lib.load().cm_adamw_step(_lib.ptr(self._p), _lib.ptr(self._g), _lib.ptr(self._m), _lib.ptr(self._v),
                                             _lib.ptr(self._numel), len(self.params), self._max_numel, C.byref(cfg), self.step_count,
                                             _lib.ptr(self._norm_ws), _lib.ptr(self._norm_out), _stream(self.flat_grad)), "cm_adamw_step")
        # the kernel wrote every parameter through its raw pointer: torch's version counters did not move, so tell the
        # caches of folded / converted weights (fusion_model._FusionRunner, RegionGraphGNN._prepared) explicitly
        _lib.bump_param_generation(self.params)
