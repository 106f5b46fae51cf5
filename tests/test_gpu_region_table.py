"""Region x ground-truth contingency table (cm_region_table_build / cm_region_table_metrics, csrc/sod_metrics.cu): the
metrics of a painted mask without a pixel pass behind the model.  Checked against (a) the oracle on the painted mask
(oracle/metrics.py = utils/metrics.py run verbatim, oracle/sod_metrics.py) and (b) the pixel kernels
(cm_paint_metrics + cm_metrics_finalize, cm_sod_metrics_painted): integer counts bit-equal, float64 sums to rounding."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _blocky_labels(rng, B, H, W, n_side=9, jitter=True):
    """SLIC-shaped label maps: a jittered grid of cells, ids 0..n-1 per image (not every id present in every image)."""
    ys, xs = np.mgrid[0:H, 0:W]
    out = np.zeros((B, H, W), np.int32)
    for b in range(B):
        wob = (rng.integers(-2, 3, size=(H, W)) if jitter else 0)
        cy = np.clip((ys + wob) * n_side // H, 0, n_side - 1)
        cx = np.clip((xs + wob) * n_side // W, 0, n_side - 1)
        out[b] = cy * n_side + cx
    return out


def _gt(rng, B, H, W, gray=False):
    ys, xs = np.mgrid[0:H, 0:W]
    g = np.zeros((B, H, W), np.uint8)
    for b in range(B):
        cy, cx, r = rng.integers(H // 4, 3 * H // 4), rng.integers(W // 4, 3 * W // 4), rng.integers(min(H, W) // 8, min(H, W) // 3)
        g[b] = (((ys - cy) ** 2 + (xs - cx) ** 2) <= r * r) * 255
    if gray:   # resize-style soft borders: bytes that are neither 0 nor 255
        soft = rng.integers(1, 255, size=g.shape).astype(np.uint8)
        edge = rng.random(g.shape) < 0.03
        g = np.where(edge, soft, g).astype(np.uint8)
    return g


def _pixel_path(labels, gt, prob, node_ptr, idm, cap, thr=0.5):
    from camouflage_multimodal_b200 import _lib, metrics
    lib = _lib.load()
    B, H, W = labels.shape
    dev = labels.device
    counts = torch.empty((B, 4), dtype=torch.int64, device=dev)
    abs_err = torch.empty(B, dtype=torch.float64, device=dev)
    per = torch.empty((B, 7), dtype=torch.float64, device=dev)
    part = torch.empty(15, dtype=torch.float64, device=dev)
    p = _lib.ptr
    paint = lib.cm_paint_metrics_i16 if labels.dtype == torch.int16 else lib.cm_paint_metrics
    _lib.check(paint(p(labels), p(gt), B, H, W, p(prob), p(node_ptr), p(idm), cap, thr, None, p(counts), p(abs_err), _lib.stream_ptr(dev)), "paint")
    _lib.check(lib.cm_metrics_finalize(p(counts), p(abs_err), B, H * W, p(per), p(part), _lib.stream_ptr(dev)), "fin")
    per_sod, part_sod = metrics.sod_metrics_painted(labels, gt, prob, node_ptr, idm, cap)
    return {"counts": counts, "abs_err": abs_err, "per_image": per, "partial": part, "per_image_sod": per_sod, "partial_sod": part_sod}


def _compare(a, b, what):
    assert torch.equal(a["counts"], b["counts"]), what
    for k, tol in (("abs_err", 1e-12), ("per_image", 1e-13), ("partial", 1e-12), ("per_image_sod", 1e-12), ("partial_sod", 1e-11)):
        x, y = a[k].cpu().numpy(), b[k].cpu().numpy()
        np.testing.assert_allclose(x, y, rtol=tol, atol=tol, equal_nan=True, err_msg=f"{what}: {k}")


def _setup(rng, B, H, W, n_side, dtype, gray=False):
    labels = _blocky_labels(rng, B, H, W, n_side)
    gt = _gt(rng, B, H, W, gray)
    n = n_side * n_side
    node_ptr = np.arange(B + 1, dtype=np.int32) * n
    prob = rng.random(B * n).astype(np.float32)
    prob[rng.random(B * n) < 0.1] = 0.0
    prob[rng.random(B * n) < 0.1] = 1.0
    dev = torch.device("cuda")
    return (torch.as_tensor(labels).to(dev, dtype), torch.as_tensor(gt).to(dev), torch.as_tensor(prob).to(dev),
            torch.as_tensor(node_ptr).to(dev), labels, gt, prob)


@pytest.mark.parametrize("dtype", [torch.int16, torch.int32])
@pytest.mark.parametrize("shape", [(4, 256, 256, 22), (3, 37, 45, 5), (2, 64, 40, 7)])
def test_table_equals_pixel_kernels_and_oracle(dtype, shape):
    from camouflage_multimodal_b200 import metrics
    from oracle import metrics as om
    from oracle import sod_metrics as osod
    B, H, W, n_side = shape
    rng = np.random.default_rng(5 + H)
    labels, gt, prob, node_ptr, labels_h, gt_h, prob_h = _setup(rng, B, H, W, n_side, dtype)
    cap = n_side * n_side + 7
    out = metrics.RegionTable(cap).build(labels, gt).metrics(prob, node_ptr)
    _compare(out, _pixel_path(labels, gt, prob, node_ptr, None, cap), f"{dtype} {shape}")
    per, per_sod = out["per_image"].cpu().numpy(), out["per_image_sod"].cpu().numpy()
    n = n_side * n_side
    for b in range(B):
        pred = prob_h[b * n:(b + 1) * n][labels_h[b]]
        ref = om.evaluate_segmentation(pred, gt_h[b] / 255.0)
        for k, name in enumerate(("iou", "dice", "precision", "recall", "f1", "mae", "accuracy")):
            tol = 1e-9 if name == "mae" else 0.0
            assert abs(per[b, k] - ref[name]) <= tol, (b, name, per[b, k], ref[name])
        sref = osod.evaluate_sod(pred.astype(np.float64), gt_h[b] / 255.0)[0]
        for k, name in enumerate(metrics.SOD_KEYS[:7]):
            assert abs(per_sod[b, k] - sref[name]) <= 1e-9, (b, name, per_sod[b, k], sref[name])
        assert np.isnan(per_sod[b, 7])


def test_table_with_non_binary_ground_truth_runs_the_pixel_kernel_for_mae():
    from camouflage_multimodal_b200 import metrics
    from oracle import metrics as om
    rng = np.random.default_rng(21)
    B, H, W, n_side = 3, 128, 128, 11
    labels, gt, prob, node_ptr, labels_h, gt_h, prob_h = _setup(rng, B, H, W, n_side, torch.int16, gray=True)
    assert ((gt_h != 0) & (gt_h != 255)).any()
    cap = 160
    out = metrics.RegionTable(cap).build(labels, gt).metrics(prob, node_ptr)
    _compare(out, _pixel_path(labels, gt, prob, node_ptr, None, cap), "gray gt")
    n = n_side * n_side
    for b in range(B):
        pred = prob_h[b * n:(b + 1) * n][labels_h[b]]
        ref = om.evaluate_segmentation(pred, gt_h[b] / 255.0)
        assert abs(out["per_image"][b, 5].item() - ref["mae"]) <= 1e-9
        assert out["per_image"][b, 0].item() == ref["iou"]


def test_table_from_bit_packed_ground_truth():
    from camouflage_multimodal_b200 import metrics
    from camouflage_multimodal_b200.pipeline import pack_mask_bits
    rng = np.random.default_rng(33)
    labels, gt, prob, node_ptr, *_ = _setup(rng, 4, 256, 256, 22, torch.int16)
    cap = 544
    packed = pack_mask_bits(gt).cuda()
    assert packed.shape == (4, 256 * 256 // 8)
    a = metrics.RegionTable(cap).build(labels, packed).metrics(prob, node_ptr)
    _compare(a, _pixel_path(labels, gt, prob, node_ptr, None, cap), "packed gt")


def test_table_label_mapping_and_out_of_range_labels():
    """id_of_label (the offset fix), raw labels >= the image's node count, negative labels and labels >= label_cap."""
    from camouflage_multimodal_b200 import metrics
    rng = np.random.default_rng(41)
    B, H, W, n_side = 2, 96, 96, 8
    labels, gt, prob, node_ptr, labels_h, *_ = _setup(rng, B, H, W, n_side, torch.int32)
    cap = 80
    lab = labels_h.copy()
    lab[0, :5, :] = -3
    lab[1, -4:, :] = cap + 100
    lab[1, 10:14, 10:30] = cap - 1          # inside the cap, beyond the node count
    labels = torch.as_tensor(lab).cuda()
    idm = torch.full((B, cap), -1, dtype=torch.int32)
    perm = rng.permutation(n_side * n_side)
    idm[:, :n_side * n_side] = torch.as_tensor(perm.astype(np.int32))
    idm = idm.cuda()
    for m, what in ((None, "raw labels"), (idm, "id_of_label")):
        out = metrics.RegionTable(cap).build(labels, gt).metrics(prob, node_ptr, m)
        _compare(out, _pixel_path(labels, gt, prob, node_ptr, m, cap), what)


def test_table_edge_cases_empty_and_full_ground_truth():
    from camouflage_multimodal_b200 import metrics
    rng = np.random.default_rng(55)
    labels, gt, prob, node_ptr, *_ = _setup(rng, 3, 64, 64, 6, torch.int16)
    gt[0] = 0
    gt[1] = 255
    out = metrics.RegionTable(40).build(labels, gt).metrics(prob, node_ptr)
    _compare(out, _pixel_path(labels, gt, prob, node_ptr, None, 40), "empty / full gt")


def test_pipeline_table_metrics_equal_pixel_metrics():
    import bench
    from camouflage_multimodal_b200 import synth
    from camouflage_multimodal_b200.pipeline import EvalPipeline, pack_mask_bits
    dev = torch.device("cuda")
    gnn, det = bench.make_models()
    kg, _ = bench.load_kg()
    gnn, det, kg = gnn.to(dev), det.to(dev), kg.to(dev)
    b = synth.make_batch(0, 8, seed=99, device=dev).as_decoded()
    outs = []
    for table, fix in ((False, False), (True, False), (False, True), (True, True)):
        pipe = EvalPipeline(gnn, det, kg, label_cap=bench.LABEL_CAP, table_metrics=table, label_offset_fix=fix)
        o = pipe.run(b.images, b.labels, b.canny, b.gt)
        torch.cuda.synchronize()
        outs.append({k: o[k].clone() for k in ("counts", "abs_err", "per_image", "partial", "per_image_sod", "partial_sod")})
    _compare(outs[1], outs[0], "pipeline")
    _compare(outs[3], outs[2], "pipeline, offset fix")
    pipe = EvalPipeline(gnn, det, kg, label_cap=bench.LABEL_CAP)
    o = pipe.run(b.images, b.labels, pack_mask_bits(b.canny).cuda(), pack_mask_bits(b.gt).cuda())
    torch.cuda.synchronize()
    _compare({k: o[k] for k in outs[0]}, outs[0], "pipeline, packed masks")
    graph, o = pipe.capture(b.images, b.labels, b.canny, b.gt)
    graph.replay()
    torch.cuda.synchronize()
    _compare({k: o[k] for k in outs[0]}, outs[0], "pipeline, captured")
