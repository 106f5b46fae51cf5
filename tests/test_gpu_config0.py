"""BASELINE configs[0]: the default model on the reference's bundled test images, batch 1, through the WHOLE path as
shipped -- SLIC -> Canny -> region graph -> RG-GNN -> RG x KG fusion + heads -> paint -- device against the CPU oracle
(oracle/slic.py, canny.py, cpu_path.py).  Inputs: tests/golden/test_images_256.npz (tools/gen_golden_test_images.py)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["img13", "img28"])
def test_bundled_image_end_to_end(name):
    from camouflage_multimodal_b200.fusion_model import build_multimodal_model
    from camouflage_multimodal_b200.region_graph import RegionGraphGNN, create_region_graph
    from oracle import canny as ocanny
    from oracle import cpu_path
    from oracle import slic as oslic
    u8 = np.load(os.path.join(os.path.dirname(__file__), "golden", "test_images_256.npz"))[name]
    image = u8.astype(np.float64) / 255.0            # the de-normalised float image of extract_rg_embeddings.py:263-276
    torch.manual_seed(0)
    gnn = RegionGraphGNN(15, 128, 2).cuda().eval()
    det = build_multimodal_model({}).cuda().eval()
    kg = torch.randn(13, 128)
    # ---- device: create_region_graph from the image alone (cm_slic + cm_canny + graph build), GNN, detector
    data, segments = create_region_graph(image, n_segments=500)
    with torch.no_grad():
        mask_nodes, _, _ = gnn(data)
        emb = gnn.extract_node_embeddings(data)
        mask_img, inst, edge, score = det(emb[None], kg[None].cuda())
    # ---- oracle: the same stages on the CPU
    seg_ref = oslic.slic((image * 255).astype(np.uint8), 500, 10.0, 1.0)
    agree = float((segments == seg_ref).mean())
    assert agree >= 0.995, f"SLIC labels agree on {agree:.4f} of the pixels"
    can_ref = ocanny.canny_of_image(image, sigma=2.0)
    gsd = {k: v.detach().cpu() for k, v in gnn.state_dict().items()}
    fsd = {k: v.detach().cpu() for k, v in det.state_dict().items()}
    gt = np.zeros((256, 256), np.uint8)
    logits_ref, _ = cpu_path.run_image(image, segments, can_ref, gt, gsd, fsd, kg, faithful=False)
    assert data.x.shape[0] == len(np.unique(segments))
    np.testing.assert_allclose(mask_img[0].cpu().numpy(), logits_ref, rtol=0, atol=1e-2)     # north star: max-abs <= 1e-2
    assert torch.isfinite(mask_nodes).all() and 0.0 <= float(score) <= 1.0
