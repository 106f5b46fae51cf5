"""The data formats either side of the hot path: the region-graph embedding cache and the RG <-> KG
matcher, so that the reference's ``train_multimodal.py`` / ``test_multimodal.py`` find the files and dicts they expect.

* ``extract_embeddings_batch`` / ``save_embedding_cache`` -- what ``extract_all_embeddings`` writes
  (models/region_graph/extract_rg_embeddings.py:360-440): ``all_rg_embeddings.pt`` =
  ``{image_name: {'node_embeddings' [N,128], 'graph_embedding' [1,128], 'num_nodes'}}``, one ``<base>_embedding.pt`` per
  image and ``embedding_summary.json``.  The embeddings come from the batched device pipeline (graph build + GNN for
  the whole batch, ONE device->host copy) instead of the reference's per-image CPU loop.
* ``EmbeddingMatcher`` -- models/multimodal/embedding_matcher.py:17-178 with the same constructor, methods, dict keys
  and matching heuristic; it also accepts in-memory dicts, and ``matched_ragged_batch`` hands the matched rows to
  ``MultimodalCamouflageDetector.forward_ragged`` without the GNN -> host -> fusion round trip of the reference
  (extract_rg_embeddings.py:288, test_multimodal.py:102).
"""
from __future__ import annotations

import json
import os

import numpy as np
import torch


def extract_embeddings_batch(gnn, images, labels, canny, names, label_cap: int | None = None):
    """Node and graph embeddings of a batch (extract_embeddings_from_image, extract_rg_embeddings.py:249-288, batched).
    ``images`` [B,H,W,3] float32 or uint8, ``labels`` [B,H,W] int32, ``canny`` [B,H,W] uint8 -- CUDA tensors.
    Returns ``{name: {'node_embeddings': [N,128] cpu, 'graph_embedding': [1,128] cpu, 'num_nodes': N}}``."""
    from .region_graph import batch_to_data, build_region_graphs
    if len(names) != images.shape[0]:
        raise ValueError("one name per image")
    g = build_region_graphs(images, labels, canny, label_cap=label_cap)
    data = batch_to_data(g)
    gnn.eval()
    with torch.no_grad():
        node = gnn.extract_node_embeddings(data)
        graph = gnn.extract_graph_embedding(data)
    ptr = g.node_ptr.cpu().tolist()
    node, graph = node.cpu(), graph.cpu()
    out = {}
    for i, name in enumerate(names):
        rows = node[ptr[i]:ptr[i + 1]].clone()
        out[name] = {"node_embeddings": rows, "graph_embedding": graph[i:i + 1].clone(), "num_nodes": rows.shape[0]}
    return out


def save_embedding_cache(all_embeddings: dict, output_dir: str, segments: dict | None = None, n_segments: int = 500):
    """Writes the files of extract_all_embeddings (:375-383, 433-460).  ``segments``: optional ``{name: label map}``."""
    os.makedirs(output_dir, exist_ok=True)
    summary = {"total_images": len(all_embeddings), "n_segments": n_segments, "embedding_dim": None, "images": {}}
    for name, rec in all_embeddings.items():
        base = os.path.splitext(name)[0]
        path = os.path.join(output_dir, f"{base}_embedding.pt")
        node, graph = rec["node_embeddings"], rec["graph_embedding"]
        one = {"image_name": name, "node_embeddings": node, "graph_embedding": graph, "num_nodes": node.shape[0],
               "embedding_dim": node.shape[1]}
        if segments is not None and name in segments:
            one["segments"] = segments[name]
        torch.save(one, path)
        summary["embedding_dim"] = int(node.shape[1])
        summary["images"][name] = {"num_nodes": int(node.shape[0]), "embedding_path": path,
                                   "node_embedding_shape": list(node.shape), "graph_embedding_shape": list(graph.shape)}
    torch.save({k: {"node_embeddings": v["node_embeddings"], "graph_embedding": v["graph_embedding"], "num_nodes": v["num_nodes"]}
                for k, v in all_embeddings.items()}, os.path.join(output_dir, "all_rg_embeddings.pt"))
    with open(os.path.join(output_dir, "embedding_summary.json"), "w") as f:
        json.dump(summary, f, indent=2)
    return summary


class EmbeddingMatcher:
    """Match Region Graph embeddings to Knowledge Graph embeddings (embedding_matcher.py:17-178)."""

    def __init__(self, rg_embeddings_path, kg_embeddings_path, category_mapping=None):
        load = lambda src: torch.load(src, weights_only=False) if isinstance(src, (str, os.PathLike)) else src  # noqa: E731
        self.rg_embeddings = load(rg_embeddings_path)
        self.kg_embeddings = load(kg_embeddings_path)
        self.category_mapping = category_mapping
        self.category_to_id = {cat: idx for idx, cat in enumerate(self.kg_embeddings.keys())}      # :45
        self.id_to_category = {idx: cat for cat, idx in self.category_to_id.items()}

    def extract_category_from_filename(self, filename):
        """'COD10K-CAM-1-Aquatic-1-Batfish-2.jpg' -> the KG category matching 'Batfish' (exact, then substring either
        way, case-insensitive), else None (:52-80)."""
        name = os.path.splitext(filename)[0]
        parts = name.split("-")
        if len(parts) >= 6:
            organism = parts[5]
            if organism in self.kg_embeddings:
                return organism
            for category in self.kg_embeddings.keys():
                if organism.lower() in category.lower() or category.lower() in organism.lower():
                    return category
        return None

    def get_kg_embedding_for_image(self, image_name, use_all_categories=False):
        """``(kg_emb [num_categories | 1, kg_dim], category_ids)`` (:82-115)."""
        if use_all_categories:
            return torch.stack(list(self.kg_embeddings.values())), list(range(len(self.kg_embeddings)))
        if self.category_mapping and image_name in self.category_mapping:
            category = self.category_mapping[image_name]
        else:
            category = self.extract_category_from_filename(image_name)
        if category and category in self.kg_embeddings:
            return self.kg_embeddings[category].unsqueeze(0), [self.category_to_id[category]]
        return torch.stack(list(self.kg_embeddings.values())).mean(dim=0, keepdim=True), [0]

    def create_matched_dataset(self, use_all_kg_categories=True):
        """List of dicts with the reference's keys (:117-165)."""
        matched = []
        for image_name, rg in self.rg_embeddings.items():
            node, graph = rg["node_embeddings"], rg["graph_embedding"]
            kg, ids = self.get_kg_embedding_for_image(image_name, use_all_categories=use_all_kg_categories)
            matched.append({"image_name": image_name, "rg_node_embeddings": node, "rg_graph_embedding": graph,
                            "kg_embeddings": kg, "category_ids": ids, "num_rg_nodes": node.shape[0],
                            "num_kg_categories": kg.shape[0]})
        return matched

    def save_matched_dataset(self, output_path, use_all_kg_categories=True):
        matched = self.create_matched_dataset(use_all_kg_categories)
        torch.save(matched, output_path)
        return matched

    def statistics(self, matched):
        """The two averages the reference prints (:158-162)."""
        return {"avg_rg_nodes": float(np.mean([d["num_rg_nodes"] for d in matched])),
                "avg_kg_categories": float(np.mean([d["num_kg_categories"] for d in matched]))}

    def matched_ragged_batch(self, image_names, device="cuda"):
        """Rows for ``MultimodalCamouflageDetector.forward_ragged``: ``(rg_rows [sum N, 128], node_ptr [B+1] int32,
        kg [1, K, kg_dim])`` with all KG categories shared by the batch (the cross-attention configuration)."""
        rows = [self.rg_embeddings[n]["node_embeddings"] for n in image_names]
        ptr = np.concatenate([[0], np.cumsum([r.shape[0] for r in rows])]).astype(np.int32)
        kg, _ = self.get_kg_embedding_for_image(image_names[0], use_all_categories=True)
        return (torch.cat(rows).to(device=device, dtype=torch.float32).contiguous(),
                torch.from_numpy(ptr).to(device), kg[None].to(device=device, dtype=torch.float32).contiguous())
