"""RG x KG fusion forward (oracle; test infrastructure only).

Functional restatement, from a state dict, of ``CrossAttentionFusion`` / ``LateFusion`` /
``MultimodalCamouflageDetector`` (models/multimodal/fusion_model.py:16-246) in eval mode, with
``nn.MultiheadAttention`` written out as explicit matmuls.  PINNED: ``oracle/make_golden.py`` runs the
reference file verbatim in the build container and ``tests/test_oracle_golden.py`` checks this module
against the committed outputs.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F


def to_3d(t: torch.Tensor, name: str) -> torch.Tensor:
    """Rank fix-ups of fusion_model.py:86-105: 2-D -> [B,1,d]; 4-D [B,a,b,d] -> squeeze a==1, else squeeze
    b==1, else view(B, a*b, d); anything else raises ValueError."""
    if t.dim() == 2:
        return t.unsqueeze(1)
    if t.dim() == 3:
        return t
    if t.dim() == 4:
        bsz, a, b, d = t.shape
        if a == 1:
            return t.squeeze(1)
        if b == 1:
            return t.squeeze(2)
        return t.reshape(bsz, a * b, d)
    raise ValueError(f"{name} must be 2D/3D/4D tensor, got shape {t.shape}")


def _drop(t, keep, pdrop):
    """F.dropout with an explicit keep mask (training parity: PyTorch's RNG stream cannot be reproduced)."""
    return t if keep is None else t * keep.to(t.dtype) / (1.0 - pdrop)


def mha(p, prefix, query, key, heads, keep=None, pdrop=0.0):
    """``nn.MultiheadAttention(batch_first=True)`` eval forward with ``need_weights=True,
    average_attn_weights=True`` (fusion_model.py:112-118): packed in-projection rows [q; k; v], q scaled by
    1/sqrt(head_dim), softmax over keys, out-projection; returns (out [B,Lq,D], head-mean probs [B,Lq,Lk])."""
    w, b = p[f"{prefix}.in_proj_weight"], p[f"{prefix}.in_proj_bias"]
    d = w.shape[1]
    hd = d // heads
    q = F.linear(query, w[:d], b[:d])
    k = F.linear(key, w[d:2 * d], b[d:2 * d])
    v = F.linear(key, w[2 * d:], b[2 * d:])
    bsz, lq, lk = query.shape[0], query.shape[1], key.shape[1]
    q = q.view(bsz, lq, heads, hd).transpose(1, 2) * (1.0 / math.sqrt(hd))
    k = k.view(bsz, lk, heads, hd).transpose(1, 2)
    v = v.view(bsz, lk, heads, hd).transpose(1, 2)
    prob = torch.softmax(q @ k.transpose(-1, -2), dim=-1)          # [B,H,Lq,Lk]
    ctx = (_drop(prob, keep, pdrop) @ v).transpose(1, 2).reshape(bsz, lq, d)
    out = F.linear(ctx, p[f"{prefix}.out_proj.weight"], p[f"{prefix}.out_proj.bias"])
    return out, prob.mean(dim=1)


def _mlp2(p, prefix, x, keep=None, pdrop=0.0):
    """Sequential(Linear, ReLU, Dropout, Linear): indices 0 and 3 (dropout only through an explicit keep mask)."""
    return F.linear(_drop(F.relu(F.linear(x, p[f"{prefix}.0.weight"], p[f"{prefix}.0.bias"])), keep, pdrop),
                    p[f"{prefix}.3.weight"], p[f"{prefix}.3.bias"])


def cross_attention_fusion(p, rg, kg, heads=8, prefix="fusion.", drop=None, pdrop=0.0):
    """fusion_model.py:75-146.  ``p`` holds the ``fusion.*`` entries of the detector state dict.  ``drop`` (training
    parity only): keep masks ``r2k`` [B,H,N,K], ``ffn_rg`` [B,N,2D], ``k2r`` [B,H,K,N], ``ffn_kg`` [B,K,2D],
    ``fusion_layer`` [B,D]."""
    dk = (lambda name: None) if drop is None else drop.get
    rg, kg = to_3d(rg, "rg_embeddings"), to_3d(kg, "kg_embeddings")
    g = lambda k: p[prefix + k]
    r = F.linear(rg, g("rg_proj.weight"), g("rg_proj.bias")) if (prefix + "rg_proj.weight") in p else rg   # :108
    k = F.linear(kg, g("kg_proj.weight"), g("kg_proj.bias")) if (prefix + "kg_proj.weight") in p else kg   # :109
    d = r.shape[-1]
    sub = {kk[len(prefix):]: v for kk, v in p.items() if kk.startswith(prefix)}
    a, w_rk = mha(sub, "cross_attn_rg2kg", r, k, heads, dk("r2k"), pdrop)                 # :112-118
    y = F.layer_norm(r + a, (d,), g("ln_rg.weight"), g("ln_rg.bias"))                    # :119
    y = y + _mlp2(sub, "ffn_rg", y, dk("ffn_rg"), pdrop)                                  # :120
    c, w_kr = mha(sub, "cross_attn_kg2rg", k, r, heads, dk("k2r"), pdrop)                 # :123-129
    z = F.layer_norm(k + c, (d,), g("ln_kg.weight"), g("ln_kg.bias"))                    # :130
    z = z + _mlp2(sub, "ffn_kg", z, dk("ffn_kg"), pdrop)                                  # :131
    both = torch.cat([y.mean(dim=1), z.mean(dim=1)], dim=-1)                              # :134-138
    fused = _mlp2(sub, "fusion_layer", both, dk("fusion_layer"), pdrop)                   # :139
    return fused, {"rg2kg": w_rk, "kg2rg": w_kr}


def late_fusion(p, rg, kg, prefix="fusion.", drop=None, pdrop=0.0):
    """fusion_model.py:164-171: mean over nodes if 3-D, concat, Linear-ReLU-Linear-ReLU-Linear (0, 3, 6).  ``drop``
    (training parity only): keep masks ``late_0`` [B,D] and ``late_1`` [B,D/2] of the two nn.Dropout sites (2, 5)."""
    dk = (lambda name: None) if drop is None else drop.get
    if rg.dim() == 3:
        rg = rg.mean(dim=1)
    if kg.dim() == 3:
        kg = kg.mean(dim=1)
    x = torch.cat([rg, kg], dim=-1)
    x = _drop(F.relu(F.linear(x, p[prefix + "fusion.0.weight"], p[prefix + "fusion.0.bias"])), dk("late_0"), pdrop)
    x = _drop(F.relu(F.linear(x, p[prefix + "fusion.3.weight"], p[prefix + "fusion.3.bias"])), dk("late_1"), pdrop)
    return F.linear(x, p[prefix + "fusion.6.weight"], p[prefix + "fusion.6.bias"]), None


def detector_forward(p, rg, kg, heads=8, fusion_type="cross_attention", drop=None, pdrop=0.0):
    """``MultimodalCamouflageDetector.forward`` (fusion_model.py:237-246), eval mode.
    Returns (mask_out, instance_out, edge_out, score_out, attention dict or None, fused)."""
    if fusion_type == "cross_attention":
        fused, attn = cross_attention_fusion(p, rg, kg, heads, drop=drop, pdrop=pdrop)
    elif fusion_type == "late":
        fused, attn = late_fusion(p, rg, kg, drop=drop, pdrop=pdrop)
    else:
        raise ValueError(f"Unknown fusion_type: {fusion_type}")
    dk = (lambda name: None) if drop is None else drop.get
    mask = _mlp2(p, "mask_head", fused, dk("mask_head"), pdrop)
    inst = _mlp2(p, "instance_head", fused, dk("instance_head"), pdrop)
    edge = _mlp2(p, "edge_head", fused, dk("edge_head"), pdrop)
    score = torch.sigmoid(_mlp2(p, "score_head", fused, dk("score_head"), pdrop))          # :229-235
    return mask, inst, edge, score, attn, fused


def fusion_flops(n: int, k: int, d_in: int = 128, d: int = 256) -> float:
    """Algorithmic FLOP per image, SURVEY.md §8(d): F_img = 2NdD + 16ND^2 + 8NKD + 10KD^2 + 10D^2."""
    return 2.0 * n * d_in * d + 16.0 * n * d * d + 8.0 * n * k * d + 10.0 * k * d * d + 10.0 * d * d
