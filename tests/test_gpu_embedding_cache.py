"""the embedding cache produced by the batched device pipeline equals per-image extraction, and feeds the
detector through the matcher without leaving the device formats the reference uses."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_batched_cache_equals_per_image_extraction_and_feeds_the_detector(tmp_path):
    from camouflage_multimodal_b200 import synth
    from camouflage_multimodal_b200.embedding_cache import EmbeddingMatcher, extract_embeddings_batch, save_embedding_cache
    from camouflage_multimodal_b200.fusion_model import build_multimodal_model
    from camouflage_multimodal_b200.region_graph import RegionGraphGNN
    torch.manual_seed(0)
    gnn = RegionGraphGNN(15, 128, 2).cuda().eval()
    b = synth.make_batch(0, 3, seed=9, device=torch.device("cuda")).as_decoded()
    labels = b.labels.to(torch.int32)
    names = [f"COD10K-CAM-1-Aquatic-1-Batfish-{i}.jpg" for i in range(3)]
    allemb = extract_embeddings_batch(gnn, b.images, labels, b.canny, names, label_cap=640)
    for i, n in enumerate(names):
        one = extract_embeddings_batch(gnn, b.images[i:i + 1], labels[i:i + 1], b.canny[i:i + 1], [n], label_cap=640)[n]
        assert one["num_nodes"] == allemb[n]["num_nodes"] == int(b.n_regions[i])
        assert float((one["node_embeddings"] - allemb[n]["node_embeddings"]).abs().max()) <= 2e-5
        assert float((one["graph_embedding"] - allemb[n]["graph_embedding"]).abs().max()) <= 2e-5
        assert allemb[n]["graph_embedding"].shape == (1, 128)
    save_embedding_cache(allemb, str(tmp_path))
    kg = {c: torch.randn(128) for c in ("Fish", "Bird", "Crab")}
    m = EmbeddingMatcher(str(tmp_path / "all_rg_embeddings.pt"), kg)
    rows, ptr, kgb = m.matched_ragged_batch(names)
    det = build_multimodal_model({}).cuda().eval()
    res = det.forward_ragged(rows, ptr, len(names), kgb)
    # the reference's calling convention on the matched dataset, sample by sample (test_multimodal.py:102)
    for i, d in enumerate(m.create_matched_dataset(use_all_kg_categories=True)):
        mask, inst, edge, score = det(d["rg_node_embeddings"][None].cuda(), d["kg_embeddings"][None].cuda())
        assert float((mask - res["mask_out"][i:i + 1]).abs().max()) <= 5e-3
        assert float((score - res["score_out"][i:i + 1]).abs().max()) <= 5e-3
