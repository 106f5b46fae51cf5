"""Canny edge map (oracle; test infrastructure only).

PARITY UNPINNED: ``skimage.feature.canny`` (scikit-image >=0.21.0, unpinned: models/region_graph/requirements.txt:4)
is neither vendored nor installed.  This restates its published algorithm on top of the SAME scipy.ndimage calls
scikit-image itself makes (``gaussian_filter``, ``sobel``, ``label``), so everything below the skimage layer is the
real third-party code; the skimage layer -- bleed-over correction, 1-pixel eroded mask, bilinear non-maximum
suppression, double threshold + connected-component hysteresis -- is written out from its documentation / source
as of 0.21-0.24.  Call site in the reference: models/region_graph/extract_rg_embeddings.py:151-152
(``gray = image . [0.2989, 0.5870, 0.1140]``; ``feature.canny(gray, sigma=2)``).
"""
from __future__ import annotations

import numpy as np
from scipy import ndimage as ndi

GRAY = np.array([0.2989, 0.5870, 0.1140])


def gray_of(image: np.ndarray) -> np.ndarray:
    """extract_rg_embeddings.py:151 written as the explicit sum (np.dot may reassociate / fuse inside BLAS)."""
    image = np.asarray(image, dtype=np.float64)
    return image[..., 0] * GRAY[0] + image[..., 1] * GRAY[1] + image[..., 2] * GRAY[2]


def smooth(gray: np.ndarray, sigma: float) -> np.ndarray:
    """skimage ``_preprocess`` with mask=None, mode='constant': Gaussian of the image divided by the Gaussian of ones."""
    mask = np.ones(gray.shape, dtype=np.float64)
    bleed_over = ndi.gaussian_filter(mask, sigma, mode="constant", cval=0.0) + np.finfo(np.float64).eps
    return ndi.gaussian_filter(gray, sigma, mode="constant", cval=0.0) / bleed_over


def nms_bilinear(isobel, jsobel, magnitude, low):
    """skimage ``_nonmaximum_suppression_bilinear``: survivors keep their magnitude, everything else is 0."""
    h, w = magnitude.shape
    out = np.zeros_like(magnitude)
    inner = np.zeros(magnitude.shape, dtype=bool)
    inner[1:-1, 1:-1] = True
    ys, xs = np.nonzero(inner & (magnitude >= low))
    for y, x in zip(ys, xs):
        m = magnitude[y, x]
        i, j = isobel[y, x], jsobel[y, x]
        down, up, left, right = i <= 0, i >= 0, j <= 0, j >= 0
        c1 = (up and right) or (down and left)
        c2 = (down and right) or (up and left)
        if not c1 and not c2:
            continue
        ai, aj = abs(i), abs(j)
        with np.errstate(divide="ignore", invalid="ignore"):
            if c1:
                if ai > aj:
                    wt = aj / ai
                    n11, n12, n21, n22 = magnitude[y + 1, x], magnitude[y + 1, x + 1], magnitude[y - 1, x], magnitude[y - 1, x - 1]
                else:
                    wt = ai / aj
                    n11, n12, n21, n22 = magnitude[y, x + 1], magnitude[y + 1, x + 1], magnitude[y, x - 1], magnitude[y - 1, x - 1]
            else:
                if ai < aj:
                    wt = ai / aj
                    n11, n12, n21, n22 = magnitude[y, x + 1], magnitude[y - 1, x + 1], magnitude[y, x - 1], magnitude[y + 1, x - 1]
                else:
                    wt = aj / ai
                    n11, n12, n21, n22 = magnitude[y - 1, x], magnitude[y - 1, x + 1], magnitude[y + 1, x], magnitude[y + 1, x - 1]
        if (n12 * wt + n11 * (1.0 - wt)) <= m and (n22 * wt + n21 * (1.0 - wt)) <= m:
            out[y, x] = m
    return out


def canny(gray: np.ndarray, sigma: float = 1.0, low_threshold: float = 0.1, high_threshold: float = 0.2) -> np.ndarray:
    """Boolean edge map of a float image (thresholds are absolute: dtype_limits of a float image is 1.0)."""
    smoothed = smooth(np.asarray(gray, dtype=np.float64), sigma)
    jsobel = ndi.sobel(smoothed, axis=1)
    isobel = ndi.sobel(smoothed, axis=0)
    magnitude = isobel * isobel
    magnitude += jsobel * jsobel
    np.sqrt(magnitude, out=magnitude)
    low_masked = nms_bilinear(isobel, jsobel, magnitude, low_threshold)
    low_mask = low_masked > 0
    labels, count = ndi.label(low_mask, np.ones((3, 3), bool))
    if count == 0:
        return low_mask
    high_mask = low_mask & (low_masked >= high_threshold)
    good = np.zeros(count + 1, dtype=bool)
    good[np.unique(labels[high_mask])] = True
    return good[labels]


def canny_of_image(image: np.ndarray, sigma: float = 2.0) -> np.ndarray:
    """extract_rg_embeddings.py:151-152."""
    return canny(gray_of(image), sigma=sigma)
