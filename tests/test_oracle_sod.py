"""oracle/sod_metrics.py (S / E / F / weighted-F; NOT in the reference -> parity unpinned): known answers of the
published definitions and a cross-check of the feature transform against scipy.  CPU only."""
import numpy as np

from oracle import sod_metrics as osod


def _gt():
    g = np.zeros((64, 80))
    g[12:40, 20:60] = 1.0
    return g


def test_perfect_and_inverted_predictions():
    g = _gt()
    m, fc, ec = osod.evaluate_sod(g, g)
    n = g.size
    assert abs(m["s_measure"] - 1.0) < 1e-12 and abs(m["wf_measure"] - 1.0) < 1e-12
    assert abs(m["e_adaptive"] - n / (n - 1)) < 1e-12 and m["f_adaptive"] == 1.0 and m["f_max"] == 1.0
    bad, _, _ = osod.evaluate_sod(1 - g, g)
    assert bad["s_measure"] == 0.0 and bad["wf_measure"] == 0.0 and bad["f_adaptive"] == 0.0 and bad["e_adaptive"] < 1e-12


def test_empty_and_full_gt_branches():
    p = np.full((16, 16), 0.25)
    assert abs(osod.s_measure(p, np.zeros((16, 16))) - 0.75) < 1e-15
    assert abs(osod.s_measure(p, np.ones((16, 16))) - 0.25) < 1e-15
    assert osod.weighted_f(p, np.zeros((16, 16))) == 0.0


def test_e_measure_closed_form_equals_the_pixel_matrix():
    """E from the four confusion counts == mean of the enhanced alignment matrix built pixel by pixel (Fan et al. eq. 3-4)."""
    rng = np.random.default_rng(0)
    g = _gt() > 0.5
    b = rng.random(g.shape) < np.where(g, 0.8, 0.1)
    tp, fp = int((b & g).sum()), int((b & ~g).sum())
    dp, dg = b - b.mean(), g - g.mean()
    align = 2 * dp * dg / (dp * dp + dg * dg + osod.EPS)
    direct = (((align + 1) ** 2) / 4).sum() / (g.size - 1 + osod.EPS)
    assert abs(float(osod.e_from_counts(np.array([tp]), np.array([fp]), int(g.sum()), g.size)[0]) - direct) < 1e-12


def test_feature_transform_distances_equal_scipy():
    from scipy import ndimage as ndi
    rng = np.random.default_rng(4)
    g = rng.random((48, 70)) < 0.02
    d2, ny, nx = osod.nearest_foreground(g)
    dst, idx = ndi.distance_transform_edt(~g, return_indices=True)
    assert np.allclose(np.sqrt(d2), dst)
    yy, xx = np.mgrid[0:48, 0:70]
    assert np.array_equal((yy - ny) ** 2 + (xx - nx) ** 2, d2) and g[ny, nx].all()      # the index really is a nearest foreground pixel
    # the tie rule only moves the weighted F by what equidistant foreground pixels differ in error
    p = np.clip(g * 0.8 + rng.random(g.shape) * 0.3, 0, 1)
    mine = osod.weighted_f(p, g.astype(float))
    e = np.abs(p - g)
    et = e.copy()
    et[~g] = e[idx[0][~g], idx[1][~g]]
    assert abs(mine - _wf_with(et, p, g, dst)) < 5e-3


def _wf_with(et, p, g, dst):
    from scipy.ndimage import convolve
    e = np.abs(p - g)
    ea = convolve(et, weights=osod.gauss7(), mode="constant", cval=0)
    m = np.where(g & (ea < e), ea, e)
    ew = m * np.where(~g, 2 - np.exp(np.log(0.5) / 5 * dst), 1.0)
    tpw, fpw = g.sum() - ew[g].sum(), ew[~g].sum()
    r, pw = 1 - ew[g].mean(), tpw / (tpw + fpw + osod.EPS)
    return 2 * r * pw / (r + pw + osod.EPS)


def test_f_curve_is_the_thresholded_f():
    rng = np.random.default_rng(1)
    g = _gt() > 0.5
    p = np.clip(g * 0.5 + rng.random(g.shape) * 0.6, 0, 1)
    fc, _ = osod.curves(p, g.astype(float))
    q = (p * 255).astype(np.uint8)
    for i in (0, 40, 128, 255):
        b = q >= 255 - i
        tp = (b & g).sum()
        pr, rc = tp / max(b.sum(), 1), tp / g.sum()
        want = 0.0 if pr + rc == 0 else 1.3 * pr * rc / (0.3 * pr + rc)
        assert abs(fc[i] - want) < 1e-12
