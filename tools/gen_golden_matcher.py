"""Golden fixture for camouflage_multimodal_b200.embedding_cache.EmbeddingMatcher from the REFERENCE's own class (build
container only: imports /root/reference/models/multimodal/embedding_matcher.py, which needs torch + numpy only).
Writes tests/golden/embedding_matcher.json: inputs (seeds, names, categories) and every matcher output in list form."""
import contextlib
import importlib.util
import io
import json
import os
import tempfile

import torch

SRC = "/root/reference/models/multimodal/embedding_matcher.py"


def reference_module():
    spec = importlib.util.spec_from_file_location("ref_embedding_matcher", SRC)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod

NAMES = ["COD10K-CAM-1-Aquatic-1-Batfish-2.jpg", "COD10K-CAM-2-Terrestrial-23-Cat-1201.jpg", "COD10K-CAM-3-Flying-53-Bird-3024.jpg",
         "COD10K-CAM-1-Aquatic-13-Pipefish-783.jpg", "plain_name.png", "COD10K-CAM-4-Amphibian-67-Frog-4500.jpg"]
CATEGORIES = ["Fish", "Cat", "Bird", "PipefishFamily", "Frog", "Crab"]
MAPPING = {"plain_name.png": "Crab"}


def inputs():
    gen = torch.Generator().manual_seed(77)
    rg = {}
    for i, n in enumerate(NAMES):
        k = 5 + 3 * i
        rg[n] = {"node_embeddings": torch.randn((k, 8), generator=gen), "graph_embedding": torch.randn((1, 8), generator=gen), "num_nodes": k}
    kg = {c: torch.randn(8, generator=gen) for c in CATEGORIES}
    return rg, kg


def main():
    mod = reference_module()
    rg, kg = inputs()
    out = {"names": NAMES, "categories": CATEGORIES, "mapping": MAPPING, "cases": {}}
    with tempfile.TemporaryDirectory() as d:
        torch.save(rg, os.path.join(d, "rg.pt"))
        torch.save(kg, os.path.join(d, "kg.pt"))
        for tag, mapping in (("no_mapping", None), ("mapping", MAPPING)):
            with contextlib.redirect_stdout(io.StringIO()):
                m = mod.EmbeddingMatcher(os.path.join(d, "rg.pt"), os.path.join(d, "kg.pt"), category_mapping=mapping)
                case = {"category_of": {n: m.extract_category_from_filename(n) for n in NAMES}, "single": {}, "all": {}}
                for n in NAMES:
                    e, ids = m.get_kg_embedding_for_image(n, use_all_categories=False)
                    case["single"][n] = {"kg": e.tolist(), "ids": ids}
                e, ids = m.get_kg_embedding_for_image(NAMES[0], use_all_categories=True)
                case["all"] = {"kg": e.tolist(), "ids": ids}
                for flag in (True, False):
                    md = m.create_matched_dataset(use_all_kg_categories=flag)
                    case[f"matched_{flag}"] = [{"image_name": r["image_name"], "category_ids": r["category_ids"], "num_rg_nodes": r["num_rg_nodes"],
                                                "num_kg_categories": r["num_kg_categories"], "kg_sum": float(r["kg_embeddings"].double().sum()),
                                                "rg_sum": float(r["rg_node_embeddings"].double().sum()), "keys": sorted(r.keys())} for r in md]
            out["cases"][tag] = case
    path = os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "embedding_matcher.json")
    json.dump(out, open(path, "w"), indent=1)
    print("wrote", path)


if __name__ == "__main__":
    main()
