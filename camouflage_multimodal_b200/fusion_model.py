"""RG x KG fusion model on the B200 kernels.

Mirrors ``models/multimodal/fusion_model.py`` of the reference: ``CrossAttentionFusion`` (:16-146),
``LateFusion`` (:149-171), ``MultimodalCamouflageDetector`` (:174-246) and ``build_multimodal_model``
(:249-259) keep their constructor arguments, forward signatures, return arities, error behaviour and
state-dict key names; the arithmetic runs through ``cm_fusion_forward``.  The parameter containers are the
stock torch modules (``nn.MultiheadAttention`` etc.) so reference checkpoints load unchanged -- their own
``forward`` is never called.
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.nn as nn

from . import _lib
from .region_graph import WORKSPACE


def _collapse_to_3d(tensor, name):
    """fusion_model.py:92-102."""
    if tensor.dim() == 3:
        return tensor
    if tensor.dim() == 4:
        b, a, bb, d = tensor.shape
        if a == 1:
            return tensor.squeeze(1)
        if bb == 1:
            return tensor.squeeze(2)
        return tensor.reshape(b, a * bb, d)
    raise ValueError(f"{name} must be 2D/3D/4D tensor, got shape {tensor.shape}")


class _FusionRunner:
    """Marshals a (fusion module, optional detector heads) pair into ``cm_fusion_params`` and runs it."""

    def __init__(self, fusion, heads=None, num_classes=2):
        self.fusion, self.heads, self.num_classes = fusion, heads, num_classes
        self._tensor_list = None
        self._key = None
        self._struct = None
        self._fold_key = None
        self._fold = None

    def _tensors(self):
        # the Parameter objects, collected once: walking the module tree costs 0.4 ms and a step asks four times
        # (modules keep their Parameter objects through .to() / load_state_dict; the key in params() sees new storage)
        if self._tensor_list is None:
            ts = list(self.fusion.parameters())
            if self.heads is not None:
                for h in self.heads:
                    ts += list(h.parameters())
            self._tensor_list = ts
        return self._tensor_list

    def params(self):
        ts = self._tensors()
        key = tuple((t.data_ptr(), t._version) for t in ts) + (_lib.param_generation(),)
        if key == self._key:
            return self._struct
        for t in ts:
            _lib.require_cuda(t, "fusion parameter")
        keep = []

        def P(t):
            t = _lib.f32c(t)
            keep.append(t)
            return t.data_ptr()

        f = self.fusion
        s = _lib.FusionParams()
        s.rg_dim, s.kg_dim, s.hidden = f.rg_dim, f.kg_dim, f.hidden_dim
        s.num_classes = self.num_classes
        if isinstance(f, CrossAttentionFusion):
            s.fusion_type, s.heads = 0, f.cross_attn_rg2kg.num_heads
            if isinstance(f.rg_proj, nn.Linear):
                s.rg_proj_w, s.rg_proj_b = P(f.rg_proj.weight), P(f.rg_proj.bias)
            if isinstance(f.kg_proj, nn.Linear):
                s.kg_proj_w, s.kg_proj_b = P(f.kg_proj.weight), P(f.kg_proj.bias)
            for pre, m in (("r2k", f.cross_attn_rg2kg), ("k2r", f.cross_attn_kg2rg)):
                setattr(s, pre + "_in_w", P(m.in_proj_weight))
                setattr(s, pre + "_in_b", P(m.in_proj_bias))
                setattr(s, pre + "_out_w", P(m.out_proj.weight))
                setattr(s, pre + "_out_b", P(m.out_proj.bias))
            s.ln_rg_w, s.ln_rg_b = P(f.ln_rg.weight), P(f.ln_rg.bias)
            s.ln_kg_w, s.ln_kg_b = P(f.ln_kg.weight), P(f.ln_kg.bias)
            s.ffn_rg_w1, s.ffn_rg_b1 = P(f.ffn_rg[0].weight), P(f.ffn_rg[0].bias)
            s.ffn_rg_w2, s.ffn_rg_b2 = P(f.ffn_rg[3].weight), P(f.ffn_rg[3].bias)
            s.ffn_kg_w1, s.ffn_kg_b1 = P(f.ffn_kg[0].weight), P(f.ffn_kg[0].bias)
            s.ffn_kg_w2, s.ffn_kg_b2 = P(f.ffn_kg[3].weight), P(f.ffn_kg[3].bias)
            s.fl_w1, s.fl_b1 = P(f.fusion_layer[0].weight), P(f.fusion_layer[0].bias)
            s.fl_w2, s.fl_b2 = P(f.fusion_layer[3].weight), P(f.fusion_layer[3].bias)
        else:
            s.fusion_type, s.heads = 1, 1
            for k, idx in enumerate((0, 3, 6)):
                s.late_w[k], s.late_b[k] = P(f.fusion[idx].weight), P(f.fusion[idx].bias)
        if self.heads is not None:
            for k, h in enumerate(self.heads):
                s.head_w1[k], s.head_b1[k] = P(h[0].weight), P(h[0].bias)
                s.head_w2[k], s.head_b2[k] = P(h[3].weight), P(h[3].bias)
        self._struct, self._key = (s, keep), key
        return self._struct

    def folded(self, kg):
        """Folded blob of ``cm_fusion_fold`` for KG rows shared by the batch (``kg`` ``[1, K, kg_dim]`` fp32), cached
        per (parameters, kg); ``None`` when the shape has no tensor-core form."""
        lib = _lib.load()
        s, _keep = self.params()
        K = kg.shape[1]
        key = (self._key, kg.data_ptr(), kg._version, K)
        if key != self._fold_key:
            nbytes = lib.cm_fusion_folded_bytes(C.byref(s), K)
            blob = None
            if nbytes:
                blob = torch.empty(nbytes, dtype=torch.uint8, device=kg.device)
                _lib.check(lib.cm_fusion_fold(C.byref(s), _lib.ptr(kg), K, _lib.ptr(blob), nbytes, _lib.stream_ptr(kg.device)),
                           "cm_fusion_fold")
            self._fold, self._fold_key = (blob, kg), key
        return self._fold[0]

    def run(self, rg_rows, node_ptr, n_graphs, kg, want_attention, want_heads, out=None):
        """rg_rows [M, rg_dim] (capacity-sized allowed), node_ptr [G+1] i32, kg [kg_graphs, K, kg_dim]."""
        lib = _lib.load()
        _lib.require_cuda(rg_rows, "rg_embeddings")
        _lib.require_cuda(kg, "kg_embeddings")
        dev = rg_rows.device
        s, _keep = self.params()
        rg_rows = _lib.f32c(rg_rows)
        if kg.dim() == 3 and kg.shape[0] > 1 and kg.stride(0) == 0:   # expand()-ed rows: one KG for the whole batch
            kg = kg[:1]
        if kg.dtype != torch.float32 or not kg.is_contiguous():
            kg = _lib.f32c(kg)
        M, G = rg_rows.shape[0], int(n_graphs)
        kg_graphs, K = kg.shape[0], kg.shape[1]
        blob = self.folded(kg) if kg_graphs == 1 else None
        final = s.hidden if s.fusion_type == 0 else s.hidden // 2
        f32 = dict(dtype=torch.float32, device=dev)
        out = out or {}
        res = {"fused": out.get("fused", torch.empty((G, final), **f32))}
        if want_heads:
            for k, c in (("mask_out", s.num_classes), ("instance_out", s.num_classes), ("edge_out", 1), ("score_out", 1)):
                res[k] = out.get(k, torch.empty((G, c), **f32))
        if want_attention and s.fusion_type == 0:
            res["rg2kg"] = out.get("rg2kg", torch.empty((M, K), **f32))
            res["kg2rg"] = out.get("kg2rg", torch.empty((K * M,), **f32))
        ws_bytes = lib.cm_fusion_workspace_bytes(C.byref(s), G, max(M, 1), K)
        if ws_bytes == 0:   # parameter validation failed: surface the library's message
            _lib.check(lib.cm_fusion_forward(C.byref(s), None, None, None, G, M, None, kg_graphs, K, None, 0, *([None] * 7),
                                             None), "cm_fusion_forward")
        ws = WORKSPACE.get(dev, "fusion", ws_bytes)
        p = _lib.ptr
        _lib.check(lib.cm_fusion_forward(C.byref(s), p(blob), p(rg_rows), p(node_ptr), G, max(M, 1), p(kg), kg_graphs, K, p(ws),
                                         ws_bytes, p(res["fused"]), p(res.get("mask_out")), p(res.get("instance_out")),
                                         p(res.get("edge_out")), p(res.get("score_out")), p(res.get("rg2kg")),
                                         p(res.get("kg2rg")), _lib.stream_ptr(dev)), "cm_fusion_forward")
        return res


def _dense_inputs(rg, kg):
    """Reference rank fix-ups (fusion_model.py:86-105) -> (rg rows [B*N, d], node_ptr, B, N, kg [Bk, K, d])."""
    if rg.dim() == 2:
        rg = rg.unsqueeze(1)
    if kg.dim() == 2:
        kg = kg.unsqueeze(1)
    rg = _collapse_to_3d(rg, "rg_embeddings")
    kg = _collapse_to_3d(kg, "kg_embeddings")
    B, N, _ = rg.shape
    if kg.shape[0] != B:
        raise RuntimeError(f"batch mismatch: rg_embeddings has {B} samples, kg_embeddings {kg.shape[0]}")
    node_ptr = torch.arange(B + 1, dtype=torch.int32, device=rg.device) * N
    return rg.reshape(B * N, rg.shape[-1]), node_ptr, B, N, kg


class CrossAttentionFusion(nn.Module):
    """fusion_model.py:16-146.  ``forward(rg, kg) -> (fused [B,hidden], {'rg2kg': [B,N,K], 'kg2rg': [B,K,N]})``."""

    def __init__(self, rg_dim=128, kg_dim=128, hidden_dim=256, num_heads=8, dropout=0.3):
        super().__init__()
        self.rg_dim, self.kg_dim, self.hidden_dim = rg_dim, kg_dim, hidden_dim
        self.dropout_p = float(dropout)
        self.rg_proj = nn.Linear(rg_dim, hidden_dim) if rg_dim != hidden_dim else nn.Identity()
        self.kg_proj = nn.Linear(kg_dim, hidden_dim) if kg_dim != hidden_dim else nn.Identity()
        self.cross_attn_rg2kg = nn.MultiheadAttention(embed_dim=hidden_dim, num_heads=num_heads, dropout=dropout, batch_first=True)
        self.cross_attn_kg2rg = nn.MultiheadAttention(embed_dim=hidden_dim, num_heads=num_heads, dropout=dropout, batch_first=True)
        self.ln_rg = nn.LayerNorm(hidden_dim)
        self.ln_kg = nn.LayerNorm(hidden_dim)
        self.ffn_rg = nn.Sequential(nn.Linear(hidden_dim, hidden_dim * 2), nn.ReLU(), nn.Dropout(dropout),
                                    nn.Linear(hidden_dim * 2, hidden_dim))
        self.ffn_kg = nn.Sequential(nn.Linear(hidden_dim, hidden_dim * 2), nn.ReLU(), nn.Dropout(dropout),
                                    nn.Linear(hidden_dim * 2, hidden_dim))
        self.fusion_layer = nn.Sequential(nn.Linear(hidden_dim * 2, hidden_dim), nn.ReLU(), nn.Dropout(dropout),
                                          nn.Linear(hidden_dim, hidden_dim))
        self._runner = _FusionRunner(self)

    def forward(self, rg_embeddings, kg_embeddings):
        rows, node_ptr, B, N, kg = _dense_inputs(rg_embeddings, kg_embeddings)
        K = kg.shape[1]
        if self.training:
            # model.train(): ``fused`` is differentiable through cm_fusion_train_forward / cm_fusion_backward (dropout
            # masks drawn here).  The attention maps are a detached side output: the head-averaged probabilities WITHOUT
            # the attention dropout nn.MultiheadAttention applies to the weights it returns in training mode (those are
            # one random draw; nothing downstream of the reference reads them while training).
            from .training import fusion_train_forward
            fused = fusion_train_forward(self, rows, node_ptr, B, kg)
            with torch.no_grad():
                res = self._runner.run(rows.detach(), node_ptr, B, kg.detach(), want_attention=True, want_heads=False)
            return fused, {"rg2kg": res["rg2kg"].view(B, N, K), "kg2rg": res["kg2rg"].view(B, K, N)}
        res = self._runner.run(rows, node_ptr, B, kg, want_attention=True, want_heads=False)
        return res["fused"], {"rg2kg": res["rg2kg"].view(B, N, K), "kg2rg": res["kg2rg"].view(B, K, N)}


class LateFusion(nn.Module):
    """fusion_model.py:149-171.  ``forward(rg, kg) -> (fused [B,hidden//2], None)``."""

    def __init__(self, rg_dim=128, kg_dim=128, hidden_dim=256, dropout=0.3):
        super().__init__()
        self.rg_dim, self.kg_dim, self.hidden_dim = rg_dim, kg_dim, hidden_dim
        self.dropout_p = float(dropout)
        self.fusion = nn.Sequential(nn.Linear(rg_dim + kg_dim, hidden_dim), nn.ReLU(), nn.Dropout(dropout),
                                    nn.Linear(hidden_dim, hidden_dim // 2), nn.ReLU(), nn.Dropout(dropout),
                                    nn.Linear(hidden_dim // 2, hidden_dim // 2))
        self._runner = _FusionRunner(self)

    @staticmethod
    def _inputs(rg, kg):
        # 3-D inputs are mean-pooled over nodes, 2-D inputs are already graph-level (:165-168)
        if rg.dim() == 2:
            rg = rg.unsqueeze(1)
        if kg.dim() == 2:
            kg = kg.unsqueeze(1)
        if rg.dim() != 3 or kg.dim() != 3:
            raise ValueError(f"LateFusion expects 2D/3D embeddings, got {tuple(rg.shape)} and {tuple(kg.shape)}")
        B, N, _ = rg.shape
        node_ptr = torch.arange(B + 1, dtype=torch.int32, device=rg.device) * N
        return rg.reshape(B * N, rg.shape[-1]), node_ptr, B, kg

    def forward(self, rg_embeddings, kg_embeddings):
        rows, node_ptr, B, kg = self._inputs(rg_embeddings, kg_embeddings)
        if self.training:   # dropout + autograd through cm_fusion_train_forward / cm_fusion_backward (fusion_type 1)
            from .training import fusion_train_forward
            return fusion_train_forward(self, rows, node_ptr, B, kg), None
        res = self._runner.run(rows, node_ptr, B, kg, want_attention=False, want_heads=False)
        return res["fused"], None


def _head(final_dim, out_dim, dropout, sigmoid=False):
    layers = [nn.Linear(final_dim, final_dim // 2), nn.ReLU(), nn.Dropout(dropout), nn.Linear(final_dim // 2, out_dim)]
    if sigmoid:
        layers.append(nn.Sigmoid())
    return nn.Sequential(*layers)


class MultimodalCamouflageDetector(nn.Module):
    """fusion_model.py:174-246: fusion + mask / instance / edge / score heads."""

    def __init__(self, rg_dim=128, kg_dim=128, hidden_dim=256, num_heads=8, fusion_type="cross_attention",
                 num_classes=2, dropout=0.3):
        super().__init__()
        self.fusion_type = fusion_type
        if fusion_type == "cross_attention":
            self.fusion = CrossAttentionFusion(rg_dim=rg_dim, kg_dim=kg_dim, hidden_dim=hidden_dim, num_heads=num_heads,
                                               dropout=dropout)
            final_dim = hidden_dim
        elif fusion_type == "late":
            self.fusion = LateFusion(rg_dim=rg_dim, kg_dim=kg_dim, hidden_dim=hidden_dim, dropout=dropout)
            final_dim = hidden_dim // 2
        else:
            raise ValueError(f"Unknown fusion_type: {fusion_type}")
        self.dropout_p = float(dropout)
        self.mask_head = _head(final_dim, num_classes, dropout)
        self.instance_head = _head(final_dim, num_classes, dropout)
        self.edge_head = _head(final_dim, 1, dropout)
        self.score_head = _head(final_dim, 1, dropout, sigmoid=True)
        self._runner = _FusionRunner(self.fusion, (self.mask_head, self.instance_head, self.edge_head, self.score_head),
                                     num_classes)

    def forward_ragged(self, rg_rows, node_ptr, n_graphs, kg, return_attention=False, out=None):
        """Ragged batch entry: all images' node embeddings as one ``[M, rg_dim]`` block, segments in ``node_ptr``
        (int32, device), ``kg`` ``[1 or n_graphs, K, kg_dim]``.  Returns the dict of ``cm_fusion_forward`` outputs."""
        if self.training:
            return self._train_forward(rg_rows, node_ptr, n_graphs, kg, return_attention)
        return self._runner.run(rg_rows, node_ptr, n_graphs, kg, want_attention=return_attention, want_heads=True, out=out)

    def _train_forward(self, rg_rows, node_ptr, n_graphs, kg, return_attention):
        """model.train(): dropout + autograd through ``cm_fusion_train_forward`` / ``cm_fusion_backward`` (training.py)."""
        from .training import fusion_train_forward
        fused, mask, inst, edge, score = fusion_train_forward(self, rg_rows, node_ptr, n_graphs, kg)
        res = {"fused": fused, "mask_out": mask, "instance_out": inst, "edge_out": edge, "score_out": score}
        if return_attention and self.fusion_type == "cross_attention":   # detached, dropout-free maps (see CrossAttentionFusion.forward)
            with torch.no_grad():
                a = self._runner.run(rg_rows.detach(), node_ptr, n_graphs, kg.detach(), want_attention=True, want_heads=False)
            res["rg2kg"], res["kg2rg"] = a["rg2kg"], a["kg2rg"]
        return res

    def forward(self, rg_embeddings, kg_embeddings, return_attention=False):
        if self.training:
            if self.fusion_type == "cross_attention":
                rows, node_ptr, B, N, kg = _dense_inputs(rg_embeddings, kg_embeddings)
            else:
                rows, node_ptr, B, kg = LateFusion._inputs(rg_embeddings, kg_embeddings)
            res = self._train_forward(rows, node_ptr, B, kg, return_attention)
            outs = (res["mask_out"], res["instance_out"], res["edge_out"], res["score_out"])
            if return_attention:
                K = kg.shape[1]
                attn = ({"rg2kg": res["rg2kg"].view(B, N, K), "kg2rg": res["kg2rg"].view(B, K, N)}
                        if self.fusion_type == "cross_attention" else None)
                return outs + (attn,)
            return outs
        attn = None
        if self.fusion_type == "cross_attention":
            rows, node_ptr, B, N, kg = _dense_inputs(rg_embeddings, kg_embeddings)
            res = self._runner.run(rows, node_ptr, B, kg, want_attention=return_attention, want_heads=True)
            if return_attention:
                K = kg.shape[1]
                attn = {"rg2kg": res["rg2kg"].view(B, N, K), "kg2rg": res["kg2rg"].view(B, K, N)}
        else:
            rows, node_ptr, B, kg = LateFusion._inputs(rg_embeddings, kg_embeddings)
            res = self._runner.run(rows, node_ptr, B, kg, want_attention=False, want_heads=True)
        outs = (res["mask_out"], res["instance_out"], res["edge_out"], res["score_out"])
        if return_attention:
            return outs + (attn,)
        return outs


def build_multimodal_model(config):
    """fusion_model.py:249-259: ``model:`` block of configs/multimodal_config.yaml -> detector."""
    return MultimodalCamouflageDetector(
        rg_dim=config.get("rg_dim", 128), kg_dim=config.get("kg_dim", 128), hidden_dim=config.get("hidden_dim", 256),
        num_heads=config.get("num_heads", 8), fusion_type=config.get("fusion_type", "cross_attention"),
        num_classes=config.get("num_classes", 2), dropout=config.get("dropout", 0.3))
