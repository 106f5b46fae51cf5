"""Region-graph GNN forward (oracle; test infrastructure only).

PARITY UNPINNED at the layer level: ``GATConv`` / ``GCNConv`` / ``global_mean_pool`` come from
torch-geometric (>=2.3.0, unpinned: models/region_graph/requirements.txt:3), which is neither vendored
nor installed.  The two conv layers below restate PyG's published message-passing semantics in plain
torch scatter ops; ``rg_gnn_forward`` then follows the reference's own ``RegionGraphGNN``
(models/region_graph/extract_rg_embeddings.py:27-135) line by line on top of them.  The layer classes
keep PyG's parameter names so that the reference class itself can be instantiated on them
(``oracle/make_golden.py`` does exactly that to cross-check ``rg_gnn_forward``).
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


def _glorot(t):
    fan = t.size(-2) + t.size(-1)
    bound = (6.0 / fan) ** 0.5
    with torch.no_grad():
        t.uniform_(-bound, bound)


def gat_conv(x, edge_index, lin_weight, att_src, att_dst, bias, heads, out_channels,
             negative_slope=0.2):
    """PyG ``GATConv(in, out, heads, concat=False)`` forward with no edge features (v1 attention).

    h = x W^T viewed [N,H,C]; a_src = <h, att_src>, a_dst = <h, att_dst> per head; self loops are removed
    then one per node is added; logit(j->i) = LeakyReLU(a_src[j] + a_dst[i]); softmax over the incoming
    edges of i (max-subtracted, denominator + 1e-16); out_i = sum_j alpha_ji h_j; mean over heads; + bias.
    edge_index[0] is the source j, edge_index[1] the target i.
    """
    n = x.shape[0]
    h = (x @ lin_weight.t()).view(n, heads, out_channels)
    a_src = (h * att_src.view(1, heads, out_channels)).sum(-1)
    a_dst = (h * att_dst.view(1, heads, out_channels)).sum(-1)
    src, dst = edge_index[0], edge_index[1]
    keep = src != dst
    loops = torch.arange(n, device=x.device, dtype=src.dtype)
    src = torch.cat([src[keep], loops])
    dst = torch.cat([dst[keep], loops])
    logit = F.leaky_relu(a_src[src] + a_dst[dst], negative_slope)           # [E', H]
    idx = dst.view(-1, 1).expand(-1, heads)
    mx = torch.full((n, heads), float("-inf"), dtype=x.dtype, device=x.device)
    mx = mx.scatter_reduce(0, idx, logit, reduce="amax", include_self=True)
    ex = (logit - mx[dst]).exp()
    den = torch.zeros((n, heads), dtype=x.dtype, device=x.device).index_add_(0, dst, ex) + 1e-16
    alpha = ex / den[dst]
    out = torch.zeros((n, heads, out_channels), dtype=x.dtype, device=x.device)
    out.index_add_(0, dst, alpha.unsqueeze(-1) * h[src])
    return out.mean(dim=1) + bias


def gcn_norm(edge_index, edge_weight, n, dtype, device):
    """PyG ``gcn_norm(add_self_loops=True, improved=False)``: existing self loops keep their weight (the
    last one per node wins), every other node gets a loop of weight 1; deg_i = sum of weights INTO i;
    w_hat(j->i) = deg_j^-1/2 * w * deg_i^-1/2 with inf -> 0."""
    src, dst = edge_index[0], edge_index[1]
    if edge_weight is None:
        edge_weight = torch.ones(src.shape[0], dtype=dtype, device=device)
    is_loop = src == dst
    loop_w = torch.ones(n, dtype=dtype, device=device)
    if bool(is_loop.any()):
        loop_w[src[is_loop]] = edge_weight[is_loop]
    ar = torch.arange(n, device=device, dtype=src.dtype)
    src = torch.cat([src[~is_loop], ar])
    dst = torch.cat([dst[~is_loop], ar])
    w = torch.cat([edge_weight[~is_loop], loop_w])
    deg = torch.zeros(n, dtype=dtype, device=device).index_add_(0, dst, w)
    dis = deg.pow(-0.5)
    dis[torch.isinf(dis)] = 0.0
    return src, dst, dis[src] * w * dis[dst]


def gcn_conv(x, edge_index, edge_weight, lin_weight, bias):
    """PyG ``GCNConv`` forward: normalise, h = x W^T (no bias in lin), scatter-add, + bias."""
    n = x.shape[0]
    src, dst, w = gcn_norm(edge_index, edge_weight, n, x.dtype, x.device)
    h = x @ lin_weight.t()
    out = torch.zeros_like(h).index_add_(0, dst, w.unsqueeze(-1) * h[src])
    return out + bias


def global_mean_pool(x, batch, num_graphs=None):
    g = int(batch.max()) + 1 if num_graphs is None else num_graphs
    s = torch.zeros((g, x.shape[1]), dtype=x.dtype, device=x.device).index_add_(0, batch, x)
    c = torch.zeros(g, dtype=x.dtype, device=x.device).index_add_(0, batch, torch.ones_like(batch, dtype=x.dtype))
    return s / c.clamp_min(1).unsqueeze(-1)


class GATConv(nn.Module):
    """Parameter container + forward with PyG >= 2.5 names (``lin.weight``, ``att_src``, ``att_dst``, ``bias``)."""

    def __init__(self, in_channels, out_channels, heads=1, concat=True):
        super().__init__()
        assert not concat, "the reference uses concat=False"
        self.heads, self.out_channels = heads, out_channels
        self.lin = nn.Linear(in_channels, heads * out_channels, bias=False)
        self.att_src = nn.Parameter(torch.empty(1, heads, out_channels))
        self.att_dst = nn.Parameter(torch.empty(1, heads, out_channels))
        self.bias = nn.Parameter(torch.zeros(out_channels))
        _glorot(self.lin.weight)
        _glorot(self.att_src)
        _glorot(self.att_dst)

    def forward(self, x, edge_index):
        return gat_conv(x, edge_index, self.lin.weight, self.att_src, self.att_dst, self.bias,
                        self.heads, self.out_channels)


class GCNConv(nn.Module):
    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.lin = nn.Linear(in_channels, out_channels, bias=False)
        self.bias = nn.Parameter(torch.zeros(out_channels))
        _glorot(self.lin.weight)

    def forward(self, x, edge_index, edge_weight=None):
        return gcn_conv(x, edge_index, edge_weight, self.lin.weight, self.bias)


def _bn_eval(x, p, name, eps=1e-5):
    return (x - p[f"{name}.running_mean"]) / torch.sqrt(p[f"{name}.running_var"] + eps) * p[f"{name}.weight"] + p[f"{name}.bias"]


def _bn_train(x, p, name, eps=1e-5):
    mu = x.mean(0)
    var = x.var(0, unbiased=False)
    return (x - mu) / torch.sqrt(var + eps) * p[f"{name}.weight"] + p[f"{name}.bias"]


def rg_gnn_forward(p, x, edge_index, edge_attr, batch=None, training_bn=False, heads=4, hidden=128, drop=None):
    """All outputs of ``RegionGraphGNN`` from a state dict ``p``.  ``drop`` (training only): list of 7 keep masks
    (or None) applied as ``F.dropout`` does -- ``h * keep / (1 - p)`` with p = 0.3 after bn1..bn3 (:61,66,71) and
    p = 0.2 after fc_shared and the three head hidden layers (:79-90); ``None`` = dropout off.

    Follows extract_rg_embeddings.py:54-135: GAT(no edge weights)+BN+ReLU, 3x GCN(edge weights)+BN+ReLU,
    fc_shared+ReLU -> node embeddings; three 2-layer heads; graph embedding = mean pool.
    ``training_bn`` uses batch statistics over all nodes (train.py:271 with Batch.from_data_list).
    """
    bn = _bn_train if training_bn else _bn_eval
    ew = edge_attr.reshape(-1) if edge_attr is not None and edge_attr.numel() > 0 else None   # :56
    lin_w = p["conv1.lin.weight"] if "conv1.lin.weight" in p else p["conv1.lin_src.weight"]
    h = gat_conv(x, edge_index, lin_w, p["conv1.att_src"], p["conv1.att_dst"], p["conv1.bias"], heads, hidden)
    def dr(t, site, pr):
        if drop is None or drop[site] is None:
            return t
        return t * drop[site].to(t.dtype) / (1.0 - pr)

    h = dr(F.relu(bn(h, p, "bn1")), 0, 0.3)
    for k in (2, 3, 4):
        h = gcn_conv(h, edge_index, ew, p[f"conv{k}.lin.weight"], p[f"conv{k}.bias"])
        h = F.relu(bn(h, p, f"bn{k}"))
        if k < 4:
            h = dr(h, k - 1, 0.3)
    emb = dr(F.relu(F.linear(h, p["fc_shared.weight"], p["fc_shared.bias"])), 3, 0.2)           # :120

    def head(n):
        t = F.relu(F.linear(emb, p[f"fc_{n}_1.weight"], p[f"fc_{n}_1.bias"]))
        t = dr(t, {"mask": 4, "instance": 5, "edge": 6}[n], 0.2)
        return F.linear(t, p[f"fc_{n}_2.weight"], p[f"fc_{n}_2.bias"])

    if batch is None:
        batch = torch.zeros(x.shape[0], dtype=torch.long)
    return {"node_emb": emb, "mask_out": head("mask"), "instance_out": head("instance"),
            "edge_out": head("edge"), "graph_emb": global_mean_pool(emb, batch) if x.shape[0] else emb.new_zeros((0, hidden))}
