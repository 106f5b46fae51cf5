"""BASELINE configs[0]: the default model on ALL 39 of the reference's bundled test images, batch 1, through the WHOLE
path as shipped -- decode bytes -> ToTensor / Normalize round trip -> SLIC -> Canny -> region graph -> RG-GNN -> RG x KG
fusion + heads -- device against the CPU oracle (oracle/slic.py, canny.py, cpu_path.py), plus the reference's own
``predict_single_image`` call sequence (models/multimodal/test_multimodal.py:83-106) on the package.
Inputs: tests/golden/test_images_256.npz (tools/gen_golden_test_images.py)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "test_images_256.npz")
NAMES = sorted(np.load(GOLDEN).files)


@pytest.fixture(scope="module")
def models():
    from camouflage_multimodal_b200.fusion_model import build_multimodal_model
    from camouflage_multimodal_b200.region_graph import RegionGraphGNN
    torch.manual_seed(0)
    gnn = RegionGraphGNN(15, 128, 2).cuda().eval()
    det = build_multimodal_model({}).cuda().eval()
    kg = torch.randn(13, 128)
    gsd = {k: v.detach().cpu() for k, v in gnn.state_dict().items()}
    fsd = {k: v.detach().cpu() for k, v in det.state_dict().items()}
    return gnn, det, kg, gsd, fsd


def _oracle(image, segments, gsd, fsd, kg):
    from oracle import canny as ocanny
    from oracle import fusion as ofus
    from oracle import gnn as ognn
    from oracle import region_graph as org
    can = ocanny.canny_of_image(image, sigma=2.0)
    x, ei, w, _ = org.build_region_graph(image, segments, can)
    with torch.no_grad():
        r = ognn.rg_gnn_forward(gsd, torch.from_numpy(x), torch.from_numpy(ei), torch.from_numpy(w))
        mask, _, _, _, _, _ = ofus.detector_forward(fsd, r["node_emb"][None], kg[None])
    return r["mask_out"].numpy(), mask[0].numpy()


def test_expected_number_of_bundled_images():
    assert len(NAMES) == 39


@pytest.mark.parametrize("name", NAMES)
def test_bundled_image_end_to_end(name, models):
    from camouflage_multimodal_b200.region_graph import create_region_graph, denormalised_image
    from oracle import slic as oslic
    gnn, det, kg, gsd, fsd = models
    u8 = np.load(GOLDEN)[name]
    image = denormalised_image(u8)                    # the float image of extract_rg_embeddings.py:263-276
    # ---- device: create_region_graph from the image alone (cm_slic + cm_canny + graph build), GNN, detector
    data, segments = create_region_graph(image, n_segments=500)
    with torch.no_grad():
        mask_nodes, _, _ = gnn(data)
        emb = gnn.extract_node_embeddings(data)
        mask_img, inst, edge, score = det(emb[None], kg[None].cuda())
    assert data.x.shape[0] == len(np.unique(segments))
    assert torch.isfinite(mask_nodes).all() and 0.0 <= float(score) <= 1.0
    # ---- oracle, first on the device's label map: isolates everything after SLIC
    nodes_ref, logits_ref = _oracle(image, segments, gsd, fsd, kg)
    np.testing.assert_allclose(mask_nodes.cpu().numpy(), nodes_ref, rtol=0, atol=1e-2)       # per-node mask logits
    np.testing.assert_allclose(mask_img[0].cpu().numpy(), logits_ref, rtol=0, atol=1e-2)     # north star: max-abs <= 1e-2
    # ---- oracle on ITS OWN SLIC output: a SLIC disagreement surfaces here
    seg_ref = oslic.slic((image * 255).astype(np.uint8), 500, 10.0, 1.0)
    agree = float((segments == seg_ref).mean())
    assert agree >= 0.995, f"SLIC labels agree on {agree:.4f} of the pixels"
    assert abs(len(np.unique(seg_ref)) - data.x.shape[0]) <= 1
    if agree < 1.0:
        _, logits_own = _oracle(image, seg_ref, gsd, fsd, kg)
        np.testing.assert_allclose(mask_img[0].cpu().numpy(), logits_own, rtol=0, atol=2e-2)


def test_predict_single_image_call_sequence(tmp_path, models):
    """The reference's predict_single_image (test_multimodal.py:83-106) on the package: extract_embeddings_from_image
    -> ordered KG tensor -> detector(rg[None], kg[None], return_attention=True), checked against the oracle."""
    from PIL import Image
    from camouflage_multimodal_b200.region_graph import extract_embeddings_from_image, load_image_256, denormalised_image
    gnn, det, kg, gsd, fsd = models
    u8 = np.load(GOLDEN)["img13"]
    path = str(tmp_path / "img13.png")               # lossless, already 256 x 256: decode + resize reproduce the bytes
    Image.fromarray(u8).save(path)
    assert np.array_equal(load_image_256(path), u8)
    device = torch.device("cuda")
    rg_node_emb, rg_graph_emb, graph_data, segments = extract_embeddings_from_image(gnn, path, device, n_segments=500)
    assert rg_node_emb.device.type == "cpu" and rg_graph_emb.shape == (1, 128) and graph_data.x.device.type == "cpu"
    assert segments.shape == (256, 256) and rg_node_emb.shape == (len(np.unique(segments)), 128)
    kg_emb = kg.unsqueeze(0).to(device)
    rg = rg_node_emb.unsqueeze(0).to(device)
    with torch.no_grad():
        outputs = det(rg, kg_emb, return_attention=True)
    assert len(outputs) == 5
    mask_out, inst_out, edge_out, score_out, attn = outputs
    assert attn["rg2kg"].shape == (1, rg.shape[1], 13) and attn["kg2rg"].shape == (1, 13, rg.shape[1])
    _, logits_ref = _oracle(denormalised_image(u8), segments, gsd, fsd, kg)
    np.testing.assert_allclose(mask_out[0].cpu().numpy(), logits_ref, rtol=0, atol=1e-2)
    torch.testing.assert_close(rg_graph_emb[0], rg_node_emb.mean(0), rtol=0, atol=1e-5)      # global_mean_pool, :133
