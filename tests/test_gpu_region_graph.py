"""GPU parity: region adjacency / compaction / features / edge weights vs the CPU oracle (A1-A4).
Integer outputs bit-exact; features abs <= 1e-5; weights rel <= 1e-5 (float32 exp of
features that themselves carry the 1e-7 feature tolerance)."""
import os

import numpy as np
import pytest
import torch

from camouflage_multimodal_b200 import synth

pytestmark = pytest.mark.gpu


def _oracle_graph(img, lab, can):
    from oracle import region_graph as org
    return org.build_region_graph(img.astype(np.float64), lab.astype(np.int64), can.astype(bool))


def _check_image(g, b, img, lab, can):
    x, ei, w, id_of = _oracle_graph(img, lab, can)
    n0, n1 = int(g.node_ptr[b]), int(g.node_ptr[b + 1])
    e0, e1 = int(g.edge_ptr[b]), int(g.edge_ptr[b + 1])
    assert n1 - n0 == x.shape[0]
    assert e1 - e0 == ei.shape[1]
    got_ei = g.edge_index[:, e0:e1].cpu().numpy() - n0           # global ids -> local
    assert np.array_equal(got_ei, ei), "edge_index differs from the canonical oracle order"
    got_id = g.id_of_label[b].cpu().numpy()
    assert np.array_equal(got_id[: len(id_of)], id_of) and (got_id[len(id_of):] == -1).all()
    np.testing.assert_allclose(g.x[n0:n1].cpu().numpy(), x, rtol=0, atol=1e-5)
    np.testing.assert_allclose(g.edge_w[e0:e1].cpu().numpy(), w.reshape(-1), rtol=1e-5, atol=1e-9)
    # CSR view is the same edge list
    rp = g.csr_row_ptr[n0:n1 + 1].cpu().numpy()
    assert rp[0] == e0 and rp[-1] == e1 and np.array_equal(np.repeat(np.arange(n1 - n0), np.diff(rp)), ei[0])
    assert np.array_equal(g.csr_col[e0:e1].cpu().numpy() - n0, ei[1])
    assert (g.batch_vec[n0:n1] == b).all()


def test_batch_matches_oracle():
    from camouflage_multimodal_b200.region_graph import build_region_graphs
    b = synth.make_batch(0, 4, seed=1234)
    d = b.to("cuda")
    g = build_region_graphs(d.images, d.labels, d.canny, label_cap=544)
    n, e = g.counts()
    assert n == int(b.n_regions.sum())
    for i in range(4):
        _check_image(g, i, b.images[i].numpy(), b.labels[i].numpy(), b.canny[i].numpy())


def test_default_label_cap_and_reuse():
    from camouflage_multimodal_b200.region_graph import build_region_graphs
    b = synth.make_batch(10, 2, seed=1234).to("cuda")
    g1 = build_region_graphs(b.images, b.labels, b.canny)
    g2 = build_region_graphs(b.images, b.labels, b.canny, label_cap=g1.label_cap, out=g1)
    assert g2 is g1
    n, e = g2.counts()
    assert n == int(b.n_regions.sum()) and e > 0


@pytest.mark.parametrize("name", ["A", "B"])
def test_golden_cases(golden_dir, name):
    """The reference's own outputs (tests/golden/region_graph.npz): case B has label 0, gaps between labels
    and a block of negative labels."""
    from camouflage_multimodal_b200.region_graph import create_region_graph
    from oracle import rag as orag
    z = np.load(os.path.join(golden_dir, "region_graph.npz"))
    spec = {"A": dict(index=3, n_regions=None), "B": dict(index=5, n_regions=23)}[name]
    sb = synth.make_batch(spec["index"], 1, seed=1234, n_regions=spec["n_regions"])
    lab = z[f"{name}.labels"].astype(np.int64)
    data, seg = create_region_graph(sb.images[0].numpy(), 500, segments=lab, edges=sb.canny[0].numpy())
    ref_ei, ref_w = orag.canonical_sort(z[f"{name}.edge_index"], z[f"{name}.edge_attr"].reshape(-1))
    assert data.edge_index.dtype == torch.int64 and data.x.dtype == torch.float32
    assert np.array_equal(data.edge_index.cpu().numpy(), ref_ei)
    np.testing.assert_allclose(data.x.cpu().numpy(), z[f"{name}.x"], rtol=0, atol=1e-5)
    assert tuple(data.edge_attr.shape) == (ref_ei.shape[1], 1)
    np.testing.assert_allclose(data.edge_attr.cpu().numpy().reshape(-1), ref_w, rtol=2e-5, atol=1e-9)
    assert seg is lab


def test_edge_cases():
    from camouflage_multimodal_b200.region_graph import build_region_graphs, create_region_graph
    # one region covering the whole frame: one node, no edges (reference returns zeros((2,0)), :237)
    img = np.random.default_rng(0).random((32, 48, 3)).astype(np.float32)
    lab = np.full((32, 48), 3, np.int32)
    can = np.zeros((32, 48), np.uint8)
    data, _ = create_region_graph(img, segments=lab, edges=can)
    assert tuple(data.x.shape) == (1, 15) and tuple(data.edge_index.shape) == (2, 0)
    assert tuple(data.edge_attr.shape) == (0, 1)
    x, ei, w, _ = _oracle_graph(img, lab, can)
    np.testing.assert_allclose(data.x.cpu().numpy(), x, atol=1e-5)
    # every pixel its own region (maximum node count), ragged non-square frame
    lab2 = np.arange(9 * 7, dtype=np.int32).reshape(9, 7)
    img2 = np.random.default_rng(1).random((9, 7, 3)).astype(np.float32)
    can2 = (np.random.default_rng(2).random((9, 7)) < 0.3).astype(np.uint8)
    data2, _ = create_region_graph(img2, segments=lab2, edges=can2)
    x2, ei2, w2, _ = _oracle_graph(img2, lab2, can2)
    assert np.array_equal(data2.edge_index.cpu().numpy(), ei2)
    np.testing.assert_allclose(data2.x.cpu().numpy(), x2, atol=1e-5)
    np.testing.assert_allclose(data2.edge_attr.cpu().numpy().reshape(-1), w2.reshape(-1), rtol=1e-5, atol=1e-9)
    # overflow is reported, not written out of bounds
    t = lambda a: torch.as_tensor(a).cuda()[None]
    g = build_region_graphs(t(img2), t(lab2), t(can2), label_cap=10)
    with pytest.raises(Exception, match="label >= label_cap"):
        g.counts()
    # all labels negative: no nodes at all
    g0 = build_region_graphs(t(img2), t(-np.ones_like(lab2)), t(can2), label_cap=4)
    assert g0.counts() == (0, 0)


def test_label_map_is_computed_when_not_given():
    """A tiny uniform image: fewer pixels than n_segments -> every pixel is its own seed; SLIC's connectivity pass then
    merges the sub-minimum segments, and the graph build runs on whatever comes out."""
    from camouflage_multimodal_b200.region_graph import create_region_graph
    data, segments = create_region_graph(np.zeros((8, 8, 3), np.float32))
    assert segments.shape == (8, 8) and data.x.shape[0] == len(np.unique(segments))


def test_full_size_properties():
    """BASELINE config 2 batch size: structural properties that need no oracle."""
    from camouflage_multimodal_b200.region_graph import build_region_graphs
    b = synth.make_batch(100, 32, seed=1234, device="cuda")
    g = build_region_graphs(b.images, b.labels, b.canny, label_cap=544)
    n, e = g.counts()
    assert n == int(b.n_regions.sum())
    ei = g.edge_index[:, :e]
    assert (ei[0] != ei[1]).all()                                   # no self loops
    key = ei[0] * n + ei[1]
    assert (key[1:] > key[:-1]).all()                               # sorted, unique
    rev = ei[1] * n + ei[0]
    assert torch.equal(torch.sort(rev).values, key)                 # symmetric
    assert torch.equal(g.batch_vec[:n], torch.repeat_interleave(torch.arange(32, device="cuda"),
                                                                torch.as_tensor(b.n_regions, device="cuda")))
    x = g.x[:n]
    assert torch.isfinite(x).all() and (x[:, 10] > 0).all()
    assert abs(float(x[:, 10].sum()) - 32.0) < 1e-3                 # region sizes tile every image
    w = g.edge_w[:e]
    assert (w > 0).all() and (w <= 1).all()
    # both directions of an edge carry the same weight
    order = torch.argsort(rev)
    assert torch.equal(w[order], w)


def test_feature_kernels_agree():
    """The pixel-parallel feature kernel (fixed-point shared-memory tables, cluster of 4 CTAs per image) and the
    bounding-box kernel (fp64 warp tree) compute the same 15 features: <= 1e-6, and each is reproducible bit for bit."""
    from camouflage_multimodal_b200.region_graph import build_region_graphs
    b = synth.make_batch(200, 8, seed=1234, device="cuda")
    g1 = build_region_graphs(b.images, b.labels, b.canny, label_cap=544)
    n, _ = g1.counts()
    x_tile = g1.x[:n].clone()
    g2 = build_region_graphs(b.images, b.labels, b.canny, label_cap=544, feature_kernel="bbox")
    x_bbox = g2.x[:n].clone()
    assert float((x_tile - x_bbox).abs().max()) <= 1e-6
    g3 = build_region_graphs(b.images, b.labels, b.canny, label_cap=544)
    assert torch.equal(g3.x[:n], x_tile)


@pytest.mark.parametrize("kernel", ["auto", "tile"])
def test_uint8_images_match_oracle_exactly(kernel):
    """Decoder-form inputs (uint8 RGB, int16 labels): the integer-moment feature kernel against the oracle run on
    byte/255.0 in float64.  The moments are exact integers, so the only difference left is the final float32
    rounding of an fp64 value that differs from numpy's in the last few ulps: tolerance 2e-7 (vs 1e-5 for float images)."""
    from camouflage_multimodal_b200.region_graph import build_region_graphs
    b = synth.make_batch(20, 3, seed=1234, n_regions=[310, 481, 530]).as_decoded()
    assert b.images.dtype == torch.uint8 and b.labels.dtype == torch.int16
    d = b.to("cuda")
    g = build_region_graphs(d.images, d.labels.to(torch.int32), d.canny, label_cap=544, feature_kernel=kernel)
    n, e = g.counts()
    assert n == int(b.n_regions.sum())
    ptr = g.node_ptr.cpu().numpy()
    for i in range(3):
        x, ei, w, _ = _oracle_graph(b.images[i].numpy().astype(np.float64) / 255.0, b.labels[i].numpy(), b.canny[i].numpy())
        got = g.x[ptr[i]:ptr[i + 1]].cpu().numpy()
        np.testing.assert_allclose(got, x, rtol=0, atol=2e-7)
        # counts-only features are bit-exact
        assert np.array_equal(got[:, 10], x[:, 10]) and np.array_equal(got[:, 13], x[:, 13]) and np.array_equal(got[:, 11], x[:, 11])


@pytest.mark.parametrize("kernel", ["auto", "tile"])
def test_uint8_odd_shapes_and_float_agreement(kernel):
    """Non-multiple-of-32 width, labels with gaps / zeros / negatives, a constant-colour image (zero variance must
    come out as exactly 0) -- uint8 path vs the float path on the same values."""
    from camouflage_multimodal_b200.region_graph import build_region_graphs
    g0 = torch.Generator().manual_seed(5)
    H, W = 72, 100
    img = torch.randint(0, 256, (2, H, W, 3), generator=g0, dtype=torch.uint8)
    img[1] = 77
    lab = (torch.arange(H)[:, None] // 9 * 12 + torch.arange(W)[None, :] // 9).to(torch.int32)[None].repeat(2, 1, 1) * 3
    lab[0, :5, :7] = -4
    lab[1, 30:40, 50:] = 0
    can = (torch.rand((2, H, W), generator=g0) < 0.1).to(torch.uint8)
    cap = int(lab.max()) + 1
    gu = build_region_graphs(img.cuda(), lab.cuda(), can.cuda(), label_cap=cap, feature_kernel=kernel)
    xu = gu.x[: gu.counts()[0]].cpu().numpy()
    gf = build_region_graphs((img.float() / 255.0).cuda(), lab.cuda(), can.cuda(), label_cap=cap)
    xf = gf.x[: gf.counts()[0]].cpu().numpy()
    assert xu.shape == xf.shape and xu.shape[0] > 50
    np.testing.assert_allclose(xu, xf, rtol=0, atol=1e-6)
    n1 = int(gu.node_ptr[1])
    assert (xu[n1:, 3:6] == 0).all() and (xu[n1:, 7] == 0).all() and (xu[n1:, 14] == 0).all() and (xu[n1:, 12] == 0).all()
    for i in range(2):
        x, _, _, _ = _oracle_graph(img[i].numpy().astype(np.float64) / 255.0, lab[i].numpy(), can[i].numpy())
        a, z = int(gu.node_ptr[i]), int(gu.node_ptr[i + 1])
        np.testing.assert_allclose(xu[a:z], x, rtol=0, atol=2e-7)


def test_labels_widen_i16():
    from camouflage_multimodal_b200 import _lib
    lib = _lib.load()
    for n in (0, 5, 8, 4099, 256 * 256 * 3 + 3):
        src = torch.randint(-32768, 32767, (max(n, 1),), dtype=torch.int16, device="cuda")[:n]
        dst = torch.full((n + 8,), 123456, dtype=torch.int32, device="cuda")
        _lib.check(lib.cm_labels_widen_i16(_lib.ptr(src) if n else _lib.ptr(dst), n, _lib.ptr(dst), _lib.stream_ptr()), "widen")
        assert torch.equal(dst[:n], src.to(torch.int32)) and (dst[n:] == 123456).all()


def test_int16_label_maps_are_read_directly_and_give_identical_graphs():
    """cm_rag_build_i16 / cm_node_features_u8_i16 on int16 labels == the int32 entry points on the widened copy."""
    from camouflage_multimodal_b200.region_graph import build_region_graphs
    b = synth.make_batch(40, 3, seed=77, device=torch.device("cuda")).as_decoded()
    assert b.labels.dtype == torch.int16 and b.images.dtype == torch.uint8
    g16 = build_region_graphs(b.images, b.labels, b.canny, label_cap=544)
    g32 = build_region_graphs(b.images, b.labels.to(torch.int32), b.canny, label_cap=544)
    n, e = g32.counts()
    assert g16.counts() == (n, e)
    for f in ("node_ptr", "edge_ptr", "id_of_label"):
        assert torch.equal(getattr(g16, f), getattr(g32, f)), f
    assert torch.equal(g16.csr_row_ptr[: n + 1], g32.csr_row_ptr[: n + 1]) and torch.equal(g16.csr_col[:e], g32.csr_col[:e])
    assert torch.equal(g16.node_bbox[:n], g32.node_bbox[:n]) and torch.equal(g16.x[:n], g32.x[:n])
    assert torch.equal(g16.edge_w[:e], g32.edge_w[:e])


def test_prefetcher_returns_the_direct_build_with_batches_in_flight():
    """RegionGraphPrefetcher (side-stream build + pinned count read): three different batches, two in flight while the
    caller's stream is busy; every graph and Batch view bit-equal to build_region_graphs + batch_to_data on the same inputs."""
    from camouflage_multimodal_b200.region_graph import RegionGraphPrefetcher, batch_to_data, build_region_graphs
    batches = [synth.make_batch(20 + 4 * k, 4, seed=77 + k, device="cuda").as_decoded() for k in range(3)]
    ins = [(b.images, b.labels.to(torch.int32), b.canny) for b in batches]
    pf = RegionGraphPrefetcher(label_cap=544)
    pf.submit(*ins[0])
    busy = torch.randn(4096, 4096, device="cuda")
    got = []
    for k in range(3):
        g, data = pf.get()
        if k + 1 < 3:
            pf.submit(*ins[k + 1])
        busy = busy @ busy * 1e-4                     # work on the caller's stream next to the side-stream build
        got.append((g, data, data.x.sum() + busy[0, 0] * 0))
    torch.cuda.synchronize()
    assert len(pf) == 0
    for (g, data, _), inp in zip(got, ins):
        ref = build_region_graphs(*inp, label_cap=544)
        rdata = batch_to_data(ref)
        n, e = ref.counts()
        assert data.x.shape[0] == n and data.edge_index.shape[1] == e and data.num_graphs == 4
        for name in ("x", "edge_w", "csr_row_ptr", "csr_col", "edge_index", "node_ptr", "edge_ptr", "id_of_label", "node_label", "node_bbox"):
            a, b = getattr(g, name), getattr(ref, name)
            lim = {"x": n, "edge_w": e, "csr_row_ptr": n + 1, "csr_col": e, "node_label": n, "node_bbox": n}.get(name)
            if name == "edge_index":
                a, b = a[:, :e], b[:, :e]
            elif lim is not None:
                a, b = a[:lim], b[:lim]
            assert torch.equal(a, b), name
        assert torch.equal(data.x, rdata.x) and torch.equal(data.edge_attr, rdata.edge_attr) and torch.equal(data.batch, rdata.batch)
        assert data._csr[0].shape[0] == n + 1 and data._csr[1].shape[0] == e


def test_prefetcher_reports_overflow_and_misuse():
    from camouflage_multimodal_b200 import _lib
    from camouflage_multimodal_b200.region_graph import RegionGraphPrefetcher
    b = synth.make_batch(0, 2, seed=5, device="cuda").as_decoded()
    with pytest.raises(ValueError):
        RegionGraphPrefetcher()
    pf = RegionGraphPrefetcher(label_cap=16)          # the label maps hold ~480 labels
    with pytest.raises(RuntimeError):
        pf.get()
    pf.submit(b.images, b.labels.to(torch.int32), b.canny)
    with pytest.raises(_lib.CmError, match="label >= label_cap"):
        pf.get()
