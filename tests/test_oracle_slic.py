"""SLIC oracle (oracle/slic.py, oracle/slic_c.c) against the REFERENCE's recorded answers: the per-image region count
its own scikit-image run produced for the 6 000 COD10K images it ships
(models/region_graph/rg_embeddings/embedding_summary.json: images[*].num_nodes).  tools/check_slic_counts.py built
tests/golden/slic_counts.json (240 images spread over the sorted list: name, expected, got) and
tests/golden/slic_images_u8.npz (the uint8 inputs of the first 64, exactly as the reference's preprocessing produces
them) in the build container.  CPU only."""
import json
import os

import numpy as np
import pytest

from oracle import slic as oslic


@pytest.fixture(scope="module")
def fixture(golden_dir):
    z = np.load(os.path.join(golden_dir, "slic_images_u8.npz"))
    return z["images"], [str(n) for n in z["names"]], z["expected"]


def test_committed_table_is_the_known_answer_check(golden_dir):
    t = json.load(open(os.path.join(golden_dir, "slic_counts.json")))
    assert t["n"] >= 200 and len(t["rows"]) == t["n"]
    exact = sum(1 for _, e, g in t["rows"] if e == g)
    assert exact == t["exact"] and exact >= 0.9 * t["n"]           # VERDICT r1 bar: >= 90 % exact, rest within 1
    assert all(abs(e - g) <= 1 for _, e, g in t["rows"])


def test_oracle_reproduces_recorded_region_counts(fixture):
    """Re-run the oracle on the committed inputs: every count equals what the reference's skimage run recorded."""
    images, names, expected = fixture
    got = np.array([len(np.unique(oslic.slic(im))) for im in images])
    assert got.tolist() == expected.tolist(), [(n, int(e), int(g)) for n, e, g in zip(names, expected, got) if e != g]


def test_table_rows_match_fixture(fixture, golden_dir):
    images, names, expected = fixture
    rows = {r[0]: r[1] for r in json.load(open(os.path.join(golden_dir, "slic_counts.json")))["rows"]}
    assert all(rows[n] == int(e) for n, e in zip(names, expected))


def test_c_restatement_equals_python_statement(fixture):
    """oracle/slic_c.c == the numpy / pure-Python loops of oracle/slic.py, bit for bit (labels, centroids)."""
    images, _, _ = fixture
    im = images[3][:96, :128]                       # a crop keeps the pure-Python walk to a second
    feat = oslic.preprocess(im)
    n_py, s_py = oslic.cluster(feat, 60, 10)
    n_c, s_c = oslic.cluster_c(feat, 60, 10)
    assert np.array_equal(n_py, n_c)
    assert np.array_equal(s_py, s_c, equal_nan=True)
    for mn, mx in ((20, 120), (60, 200), (1, 9)):
        assert np.array_equal(oslic.enforce_connectivity(n_py + 1, mn, mx, 1), oslic.enforce_connectivity_c(n_py + 1, mn, mx, 1))
    assert np.array_equal(oslic.slic(im, 60, use_c=False), oslic.slic(im, 60, use_c=True))


def test_small_first_segment_joins_the_first_label():
    """The rule the recorded counts pinned (oracle/slic.py enforce_connectivity docstring): no label start_label - 1."""
    lab = np.ones((8, 8), np.int64)
    lab[0, 0] = 7                                    # a 1-pixel segment met first by the raster scan
    for fn in (oslic.enforce_connectivity, oslic.enforce_connectivity_c):
        out = fn(lab, 4, 100, 1)
        assert out.min() == 1 and (out == 1).all()


def test_preprocess_order_and_rescale():
    """skimage >= 0.19: min-max rescale, Lab first, blur after.  A constant image stays finite; a low-contrast image
    is stretched to the full range before Lab."""
    const = np.full((16, 16, 3), 77, np.uint8)
    assert np.isfinite(oslic.preprocess(const)).all()
    rng = np.random.default_rng(0)
    lo = (rng.integers(0, 52, size=(32, 32, 3)) + 50).astype(np.uint8)      # range 51 = 255 / 5: the stretch is exact in bytes
    lo[0, 0, 0], lo[0, 0, 1] = 50, 101
    stretched = ((lo.astype(np.int64) - 50) * 5).astype(np.uint8)
    a, b = oslic.preprocess(lo), oslic.preprocess(stretched)
    assert np.abs(a - b).max() < 1e-9


@pytest.mark.skipif(not os.path.isdir("/root/reference/data/COD10K/images"), reason="the reference tree only exists in the build container")
def test_oracle_counts_straight_from_the_reference_files():
    """Build container only: a different sample than the committed table (every 251st image from an offset), decoded and
    preprocessed from the reference's own files, against embedding_summary.json."""
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tools"))
    from check_slic_counts import REF, reference_uint8
    summ = json.load(open(f"{REF}/models/region_graph/rg_embeddings/embedding_summary.json"))["images"]
    names = sorted(summ)[17::251]
    bad = []
    for n in names:
        _, u8 = reference_uint8(f"{REF}/data/COD10K/images/{n}")
        got = len(np.unique(oslic.slic(u8)))
        if got != summ[n]["num_nodes"]:
            bad.append((n, summ[n]["num_nodes"], got))
    assert len(names) >= 20 and len(bad) <= len(names) // 10 and all(abs(e - g) <= 1 for _, e, g in bad), bad
