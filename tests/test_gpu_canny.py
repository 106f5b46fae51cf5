"""GPU parity of the Canny front end against oracle/canny.py (scipy.ndimage underneath, skimage's
layer restated; parity unpinned).  Everything runs in float64 in scipy's operation order on both sides, so the
boolean maps are expected to be IDENTICAL; the assertion allows 0.02 % of the pixels for floating-point ties on the
non-maximum-suppression comparisons and on the thresholds."""
import numpy as np
import pytest
import torch

from camouflage_multimodal_b200 import synth

pytestmark = pytest.mark.gpu


def _structured_image(seed, h=256, w=256):
    """Smooth blobs + a few sharp shapes, so that there are real edges for sigma = 2."""
    g = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    img = np.zeros((h, w, 3))
    for _ in range(6):
        cy, cx, r = g.uniform(0, h), g.uniform(0, w), g.uniform(12, 60)
        col = g.uniform(0.2, 1.0, 3)
        img[(yy - cy) ** 2 + (xx - cx) ** 2 < r * r] = col
    img[40:90, 150:230] = g.uniform(0, 1, 3)
    img += g.normal(0, 0.02, img.shape)
    return np.clip(img, 0, 1).astype(np.float32)


def _mismatch(a, b):
    return float((a != b).mean())


def test_canny_matches_oracle_float_and_uint8():
    from camouflage_multimodal_b200.region_graph import canny_edges
    from oracle import canny as oc
    imgs = np.stack([_structured_image(s) for s in (1, 2, 3)])
    got = canny_edges(torch.from_numpy(imgs).cuda(), sigma=2.0).cpu().numpy().astype(bool)
    for i in range(3):
        ref = oc.canny_of_image(imgs[i].astype(np.float64), sigma=2.0)
        assert ref.sum() > 500, "test image should have edges"
        assert _mismatch(got[i], ref) <= 2e-4, (got[i].sum(), ref.sum())
        assert not got[i][0].any() and not got[i][-1].any() and not got[i][:, 0].any() and not got[i][:, -1].any()
    u8 = (imgs * 255).round().astype(np.uint8)
    got8 = canny_edges(torch.from_numpy(u8).cuda(), sigma=2.0).cpu().numpy().astype(bool)
    for i in range(3):
        ref = oc.canny_of_image(u8[i].astype(np.float64) / 255.0, sigma=2.0)
        assert _mismatch(got8[i], ref) <= 2e-4


@pytest.mark.parametrize("shape,sigma", [((67, 131), 1.0), ((300, 200), 3.0), ((256, 256), 2.0)])
def test_canny_shapes_sigmas_and_noise(shape, sigma):
    """Odd shapes (border handling of both filters), other sigmas, and the bench's low-contrast synthetic images."""
    from camouflage_multimodal_b200.region_graph import canny_edges
    from oracle import canny as oc
    h, w = shape
    img = _structured_image(7, h, w) if shape != (256, 256) else synth.smooth_image(1234, 5).numpy()
    got = canny_edges(torch.from_numpy(img).cuda()[None], sigma=sigma)[0].cpu().numpy().astype(bool)
    ref = oc.canny_of_image(img.astype(np.float64), sigma=sigma)
    assert _mismatch(got, ref) <= 2e-4, (got.sum(), ref.sum())


def test_create_region_graph_computes_canny_when_not_given():
    """Feature 13 (mean of the Canny map over the region, :182) from the device-side Canny equals the one from the
    oracle's map handed in explicitly."""
    from camouflage_multimodal_b200.region_graph import create_region_graph
    from oracle import canny as oc
    img = _structured_image(11)
    lab = synth.voronoi_labels(300, 1234, 3).numpy().astype(np.int64)
    d1, _ = create_region_graph(img, 500, segments=lab)
    d2, _ = create_region_graph(img, 500, segments=lab, edges=oc.canny_of_image(img.astype(np.float64), 2.0))
    assert torch.equal(d1.edge_index, d2.edge_index)
    assert (d1.x[:, 13] - d2.x[:, 13]).abs().max() <= 0.05      # a tie pixel moves a region's mean by 1/area
    assert (d1.x[:, 13] != d2.x[:, 13]).float().mean() <= 0.02
    assert torch.equal(d1.x[:, :13], d2.x[:, :13])
