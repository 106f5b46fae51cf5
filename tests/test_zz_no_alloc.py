"""cm_sod_metrics_static: the S- / E- / F-measure reductions with their scratch in a device array built into the library
(no allocation in the call; include/cm_b200.h).  No counterpart in the reference.
Called through the C ABI and compared with cm_sod_metrics (caller workspace, with_weighted_f = 0) on the same
inputs, two calls in a row with different batches: the second finds the first call's contents in the shared array.
Collected last on purpose (the file name)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
TOL = 1e-9   # the tolerance of tests/test_gpu_sod_metrics.py; both paths run the same kernels, so they agree to the bit


def _inputs(seed, B, H, W):
    rng = np.random.default_rng(seed)
    ys, xs = np.mgrid[0:H, 0:W]
    gt = np.zeros((B, H, W), np.uint8)
    pred = np.zeros((B, H, W), np.float32)
    for b in range(B):
        cy, cx, r = rng.integers(H // 4, 3 * H // 4), rng.integers(W // 4, 3 * W // 4), rng.integers(min(H, W) // 8, min(H, W) // 3)
        g = ((ys - cy) ** 2 + (xs - cx) ** 2) <= r * r
        gt[b] = g * 255
        pred[b] = np.clip(np.roll(g, int(rng.integers(-6, 7)), axis=1) * 0.8 + rng.random((H, W)) * 0.3, 0, 1)
    dev = torch.device("cuda")
    return torch.as_tensor(pred).to(dev), torch.as_tensor(gt).to(dev)


def _reference_entry(lib, _lib, pred, gt):
    B, H, W = pred.shape
    dev = pred.device
    per = torch.empty((B, 8), dtype=torch.float64, device=dev)
    part = torch.empty(528, dtype=torch.float64, device=dev)
    ws_bytes = lib.cm_sod_workspace_bytes(B, H, W)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    p = _lib.ptr
    _lib.check(lib.cm_sod_metrics(p(pred), p(gt), B, H, W, 0, p(ws), ws_bytes, p(per), p(part), _lib.stream_ptr(dev)), "cm_sod_metrics")
    return per, part


def _static_entry(lib, _lib, pred, gt):
    B, H, W = pred.shape
    dev = pred.device
    per = torch.full((B, 8), -7.0, dtype=torch.float64, device=dev)
    part = torch.full((528,), -7.0, dtype=torch.float64, device=dev)
    p = _lib.ptr
    rc = lib.cm_sod_metrics_static(p(pred), p(gt), B, H, W, p(per), p(part), _lib.stream_ptr(dev))
    return rc, per, part


def test_static_scratch_entry_matches_workspace_entry_twice_in_a_row():
    from camouflage_multimodal_b200 import _lib
    lib = _lib.load()
    assert lib.cm_sod_metrics_static_max_bytes() == 2 << 20
    for seed, (B, H, W) in ((1, (6, 256, 256)), (2, (3, 96, 130))):   # the second call: fewer, smaller images over stale scratch
        pred, gt = _inputs(seed, B, H, W)
        rc, per, part = _static_entry(lib, _lib, pred, gt)
        _lib.check(rc, "cm_sod_metrics_static")
        ref_per, ref_part = _reference_entry(lib, _lib, pred, gt)
        torch.cuda.synchronize()
        a, b = per.cpu().numpy(), ref_per.cpu().numpy()
        assert np.isnan(a[:, 7]).all() and np.isnan(b[:, 7]).all()        # weighted F is not part of either call
        np.testing.assert_allclose(a[:, :7], b[:, :7], rtol=0, atol=TOL)
        np.testing.assert_allclose(part.cpu().numpy(), ref_part.cpu().numpy(), rtol=0, atol=TOL, equal_nan=True)
        assert np.isfinite(a[:, :7]).all() and (a[:, :7] > -1e-12).all() and a[:, 0].max() > 0.0   # written, and not trivially zero


def test_static_scratch_entry_rejects_batches_beyond_the_array():
    from camouflage_multimodal_b200 import _lib
    lib = _lib.load()
    B, H, W = 400, 256, 256          # ~12 KB of scratch per 256 x 256 image: 400 images do not fit 2 MiB
    dev = torch.device("cuda")
    pred = torch.zeros((B, H, W), dtype=torch.float32, device=dev)
    gt = torch.zeros((B, H, W), dtype=torch.uint8, device=dev)
    rc, per, _ = _static_entry(lib, _lib, pred, gt)
    assert rc == _lib.CM_ERR_WORKSPACE
    assert "built-in array" in lib.cm_last_error().decode()
    torch.cuda.synchronize()
    assert (per == -7.0).all()       # nothing ran
