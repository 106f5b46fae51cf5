"""Minimal stand-ins for ``torch_geometric.data.Data`` / ``Batch``.

The reference builds one ``Data(x, edge_index, edge_attr[, y, instance_y, edge_y])`` per image
(models/region_graph/extract_rg_embeddings.py:240-244, train.py:203-210) and joins them with
``Batch.from_data_list`` (train.py:231): a disjoint union whose node ids are offset per graph plus a
``batch`` vector.  PyTorch-Geometric is not a dependency here, so the same duck type is provided.
"""
from __future__ import annotations

import torch


class Data:
    def __init__(self, x=None, edge_index=None, edge_attr=None, **extra):
        self.x = x
        self.edge_index = edge_index
        self.edge_attr = edge_attr
        for k, v in extra.items():
            setattr(self, k, v)

    def keys(self):
        return [k for k, v in self.__dict__.items() if v is not None and not k.startswith("_")]

    def _apply(self, fn):
        out = self.__class__.__new__(self.__class__)
        for k, v in self.__dict__.items():
            out.__dict__[k] = fn(v) if torch.is_tensor(v) else v
        return out

    def to(self, device, non_blocking=False):
        return self._apply(lambda t: t.to(device, non_blocking=non_blocking))

    def cpu(self):
        return self._apply(lambda t: t.cpu())

    def cuda(self, device=None):
        return self._apply(lambda t: t.cuda(device))

    @property
    def num_nodes(self):
        return 0 if self.x is None else self.x.shape[0]

    @property
    def num_edges(self):
        return 0 if self.edge_index is None else self.edge_index.shape[1]

    def __repr__(self):
        parts = [f"{k}={list(getattr(self, k).shape)}" for k in self.keys() if torch.is_tensor(getattr(self, k))]
        return f"{self.__class__.__name__}({', '.join(parts)})"


class Batch(Data):
    @classmethod
    def from_data_list(cls, data_list):
        if len(data_list) == 0:
            raise ValueError("empty data list")
        xs, eis, eas, bvec, ptr = [], [], [], [], [0]
        extras = {}
        for g, d in enumerate(data_list):
            n = d.x.shape[0]
            xs.append(d.x)
            eis.append(d.edge_index + ptr[-1])
            eas.append(d.edge_attr if d.edge_attr is not None else d.x.new_zeros((0, 1)))
            bvec.append(torch.full((n,), g, dtype=torch.long, device=d.x.device))
            ptr.append(ptr[-1] + n)
            for k in d.keys():
                if k not in ("x", "edge_index", "edge_attr"):
                    extras.setdefault(k, []).append(getattr(d, k))
        out = cls(x=torch.cat(xs, 0), edge_index=torch.cat(eis, 1), edge_attr=torch.cat(eas, 0))
        out.batch = torch.cat(bvec, 0)
        out.ptr = torch.tensor(ptr, dtype=torch.long, device=out.x.device)
        out.num_graphs = len(data_list)
        for k, vs in extras.items():
            if all(torch.is_tensor(v) for v in vs):
                setattr(out, k, torch.cat([v.reshape(-1) if v.dim() == 0 else v for v in vs], 0))
        return out
