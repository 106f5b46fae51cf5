"""ORACLE (test infrastructure only -- never imported by the product): CPU restatement of the training-loop pieces
around the model passes, following the reference line by line.

* region_targets      models/region_graph/train.py:169-176 (masks are ToTensor'ed: float32 = byte / 255)
* rg_node_loss        models/region_graph/train.py:325-353
* AggressiveFocalLoss models/multimodal/train_multimodal.py:29-57
* multimodal_loss     models/multimodal/train_multimodal.py:253-268 (one backward per sample -> sum over the batch)
* clip + AdamW        torch.nn.utils.clip_grad_norm_ / torch.optim.AdamW themselves (train.py:318, 356-357): the
                      reference calls the library, so the library on the CPU is the oracle.
"""
import numpy as np
import torch
import torch.nn.functional as F


def region_targets(labels, mask_u8, inst_u8, edge_u8):
    """One image.  labels [H,W] int (regions in ascending label order, as the reference's np.unique loop visits them);
    masks uint8.  Returns (y, instance_y, edge_y) lists."""
    mask = mask_u8.astype(np.float32) / np.float32(255)
    inst = inst_u8.astype(np.float32) / np.float32(255)
    edge = edge_u8.astype(np.float32) / np.float32(255)
    y, iy, ey = [], [], []
    for rid in np.unique(labels):
        m = labels == rid
        y.append(1 if mask[m].mean() > 0.5 else 0)          # :170
        iy.append(1 if inst[m].mean() > 0.5 else 0)         # :173
        ey.append(1 if edge[m].mean() > 0.3 else 0)         # :176
    return y, iy, ey


def rg_node_loss(mask_out, inst_out, edge_out, y, iy, ey):
    """train.py:325-353 in the dtype of the inputs (float64 in the tests)."""
    dt = mask_out.dtype
    lm = F.cross_entropy(mask_out, y, weight=torch.tensor([1.0, 5.0], dtype=dt)) * 2.0
    li = F.cross_entropy(inst_out, iy, weight=torch.tensor([1.0, 4.0], dtype=dt)) * 1.0
    le = F.binary_cross_entropy_with_logits(edge_out.reshape(-1), ey.to(dt), pos_weight=torch.tensor([3.0], dtype=dt)) * 0.5
    return lm + li + le, lm, li, le


def aggressive_focal_loss(inputs, targets, alpha=0.75, gamma=3.0):
    ce = F.cross_entropy(inputs, targets, reduction="none")                     # :45
    probs = F.softmax(inputs, dim=1)
    pt = probs.gather(1, targets.unsqueeze(1)).squeeze(1)                        # :47
    alpha_t = torch.where(targets == 1, torch.tensor(alpha, dtype=inputs.dtype), torch.tensor(1 - alpha, dtype=inputs.dtype))
    return (alpha_t * (1 - pt) ** gamma * ce).mean()                             # :53-56


def multimodal_loss(mask_out, inst_out, edge_out, score_out, mask_label, edge_label, score_label):
    """Sum over samples of the per-sample loss of train_multimodal.py:253-268 (batch of one per forward)."""
    total = 0.0
    for g in range(mask_out.shape[0]):
        ml = mask_label[g:g + 1]
        lm = aggressive_focal_loss(mask_out[g:g + 1], ml) * 3.0
        li = F.cross_entropy(inst_out[g:g + 1], ml) * 1.0
        le = F.binary_cross_entropy_with_logits(edge_out[g:g + 1].reshape(-1), edge_label[g:g + 1].to(mask_out.dtype)) * 0.5
        ls = F.mse_loss(score_out[g:g + 1].reshape(-1), score_label[g:g + 1].to(mask_out.dtype)) * 0.3
        total = total + lm + li + le + ls
    return total
