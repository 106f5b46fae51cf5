"""Seeded synthetic COD10K-shaped inputs for the region-graph hot path.

SURVEY.md §8(d): images are 256x256 RGB f32 in [0,1]; superpixel label maps are
jittered-grid Voronoi partitions whose region counts follow the empirical
``num_nodes`` histogram of the reference's 6 000-image extraction log
(models/region_graph/rg_embeddings/embedding_summary.json: min 303, mean 480.9,
max 530); the Canny map is Bernoulli(0.08); ground truth is 1-3 filled ellipses.

Every image ``i`` of a set is a pure function of ``(seed, i)`` so that any subset
(a rank's shard, the CPU-baseline sample) sees exactly the images the full set has.
All integer outputs (labels, canny, gt) are computed with integer arithmetic only and
are therefore identical whichever torch device generates them.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np
import torch

H = 256
W = 256

# 10-wide bins [300,310) ... [530,540) of num_nodes over the reference's 6 000 images.
_HIST_LO = 300
_HIST_COUNTS = np.array(
    [2, 0, 1, 3, 3, 2, 7, 11, 28, 27, 40, 69, 113, 169, 289, 406, 568, 713, 896, 996, 877, 592, 187, 1],
    dtype=np.float64,
)
_N_MIN, _N_MAX = 303, 530


def draw_region_counts(n_images: int, seed: int = 1234, first: int = 0) -> np.ndarray:
    """Region count per image, images ``first .. first+n_images-1`` of the seeded set."""
    out = np.empty(n_images, dtype=np.int64)
    p = _HIST_COUNTS / _HIST_COUNTS.sum()
    for k in range(n_images):
        rng = np.random.default_rng([seed, first + k, 17])
        b = rng.choice(len(p), p=p)
        out[k] = int(np.clip(_HIST_LO + 10 * b + rng.integers(0, 10), _N_MIN, _N_MAX))
    return out


def _hash32(*terms: torch.Tensor) -> torch.Tensor:
    """Integer mixing hash on int64 tensors, result in [0, 2^32)."""
    m = 0xFFFFFFFF
    h = torch.zeros_like(terms[0])
    for k, t in enumerate(terms):
        h = (h ^ ((t + (0x9E3779B9 + 0x85EBCA6B * k)) & m)) & m
        h = (h * 0x01000193) & m
        h = (h ^ (h >> 15)) & m
        h = (h * 0x2C1B3C6D) & m
        h = (h ^ (h >> 12)) & m
    return h


def voronoi_labels(n_regions: int, seed: int, index: int, device="cpu", h: int = H, w: int = W) -> torch.Tensor:
    """One ``[h,w]`` int32 label map with labels ``1..n_regions`` (every label present)."""
    n_regions = int(n_regions)
    if not (1 <= n_regions <= h * w):
        raise ValueError(f"n_regions={n_regions} does not fit a {h}x{w} image")
    gx = max(1, int(math.ceil(math.sqrt(n_regions * w / h))))
    gy = max(1, int(math.ceil(n_regions / gx)))
    while gx > w or gy > h:  # cells must be at least one pixel
        gx = min(gx, w)
        gy = int(math.ceil(n_regions / gx))
        if gy > h:
            raise ValueError("too many regions")
    rng = np.random.default_rng([seed, index, 29])
    keep = np.ones(gy * gx, dtype=bool)
    n_drop = gy * gx - n_regions
    if n_drop:
        keep[rng.choice(gy * gx, size=n_drop, replace=False)] = False
    # cell c covers rows [ry[c], ry[c+1]) ; the seed is a jittered interior pixel
    ry = (np.arange(gy + 1) * h) // gy
    rx = (np.arange(gx + 1) * w) // gx
    u = rng.uniform(0.15, 0.85, size=(gy, gx))
    v = rng.uniform(0.15, 0.85, size=(gy, gx))
    sy = (ry[:-1, None] + np.floor(u * (ry[1:] - ry[:-1])[:, None])).astype(np.int64)
    sx = (rx[None, :-1] + np.floor(v * (rx[1:] - rx[:-1])[None, :])).astype(np.int64)
    lab = np.zeros(gy * gx, dtype=np.int64)
    lab[keep] = np.arange(1, n_regions + 1)
    lab = lab.reshape(gy, gx)

    dev = torch.device(device)
    sy_t = torch.from_numpy(sy).to(dev)
    sx_t = torch.from_numpy(sx).to(dev)
    lab_t = torch.from_numpy(lab).to(dev)
    py = torch.arange(h, device=dev)
    px = torch.arange(w, device=dev)
    cy = torch.bucketize(py, torch.from_numpy(ry[1:-1].copy()).to(dev), right=True)  # [h] cell row
    cx = torch.bucketize(px, torch.from_numpy(rx[1:-1].copy()).to(dev), right=True)  # [w]
    big = torch.iinfo(torch.int64).max // 4
    best_d = torch.full((h, w), big, dtype=torch.int64, device=dev)
    best_l = torch.zeros((h, w), dtype=torch.int64, device=dev)
    R = 2
    for dy in range(-R, R + 1):
        ny = cy + dy
        oky = (ny >= 0) & (ny < gy)
        nyc = ny.clamp(0, gy - 1)
        for dx in range(-R, R + 1):
            nx = cx + dx
            okx = (nx >= 0) & (nx < gx)
            nxc = nx.clamp(0, gx - 1)
            cand_y = sy_t[nyc][:, nxc]
            cand_x = sx_t[nyc][:, nxc]
            cand_l = lab_t[nyc][:, nxc]
            d = (py[:, None] - cand_y) ** 2 + (px[None, :] - cand_x) ** 2
            ok = oky[:, None] & okx[None, :] & (cand_l > 0)
            d = torch.where(ok, d, torch.full_like(d, big))
            better = d < best_d
            best_d = torch.where(better, d, best_d)
            best_l = torch.where(better, cand_l, best_l)
    if bool((best_l == 0).any()):
        # A 5x5 cell window held no kept seed (only when many cells were dropped): fall back to all seeds.
        ys, xs = np.nonzero(lab)
        all_y = torch.from_numpy(sy[ys, xs]).to(dev)
        all_x = torch.from_numpy(sx[ys, xs]).to(dev)
        all_l = torch.from_numpy(lab[ys, xs]).to(dev)
        miss = (best_l == 0).nonzero()
        d = (miss[:, 0:1] - all_y[None]) ** 2 + (miss[:, 1:2] - all_x[None]) ** 2
        best_l[miss[:, 0], miss[:, 1]] = all_l[d.argmin(dim=1)]
    return best_l.to(torch.int32)


def smooth_image(seed: int, index: int, device="cpu", h: int = H, w: int = W) -> torch.Tensor:
    """``[h,w,3]`` f32 in [0,1]: bilinear-upsampled coarse noise plus fine hash noise."""
    g = torch.Generator().manual_seed((seed * 1000003 + index) & 0x7FFFFFFF)
    coarse = torch.rand((1, 3, 34, 34), generator=g).to(device)
    img = torch.nn.functional.interpolate(coarse, size=(h, w), mode="bilinear", align_corners=False)[0]
    yy = torch.arange(h, device=device, dtype=torch.int64)[:, None].expand(h, w)
    xx = torch.arange(w, device=device, dtype=torch.int64)[None, :].expand(h, w)
    idx = torch.full_like(yy, index)
    fine = torch.stack([_hash32(yy, xx, idx, torch.full_like(yy, 101 + c)) for c in range(3)], 0)
    fine = (fine & 0xFFFF).to(torch.float32) / 65535.0 - 0.5
    img = (img * 0.84 + 0.08 + 0.16 * fine).clamp_(0.0, 1.0)
    return img.permute(1, 2, 0).contiguous()


def canny_like(seed: int, index: int, device="cpu", h: int = H, w: int = W, p: float = 0.08) -> torch.Tensor:
    """``[h,w]`` u8 in {0,1}: Bernoulli(p) stand-in for skimage.feature.canny."""
    yy = torch.arange(h, device=device, dtype=torch.int64)[:, None].expand(h, w)
    xx = torch.arange(w, device=device, dtype=torch.int64)[None, :].expand(h, w)
    r = _hash32(yy, xx, torch.full_like(yy, index), torch.full_like(yy, seed & 0xFFFF)) & 0xFFFF
    return (r < int(p * 65536)).to(torch.uint8)


def ellipse_gt(seed: int, index: int, device="cpu", h: int = H, w: int = W) -> torch.Tensor:
    """``[h,w]`` u8 in {0,255}: union of 1-3 filled ellipses (COD10K gt_object style)."""
    rng = np.random.default_rng([seed, index, 43])
    yy = torch.arange(h, device=device, dtype=torch.float32)[:, None]
    xx = torch.arange(w, device=device, dtype=torch.float32)[None, :]
    m = torch.zeros((h, w), dtype=torch.bool, device=device)
    for _ in range(int(rng.integers(1, 4))):
        cy, cx = rng.uniform(0.2 * h, 0.8 * h), rng.uniform(0.2 * w, 0.8 * w)
        a, b = rng.uniform(0.06 * w, 0.3 * w), rng.uniform(0.06 * h, 0.3 * h)
        th = rng.uniform(0, math.pi)
        # integer-valued coordinates and f32 products: the mask may differ between devices only
        # on exact-boundary pixels, which have measure zero for these irrational parameters.
        u = (xx - cx) * math.cos(th) + (yy - cy) * math.sin(th)
        v = -(xx - cx) * math.sin(th) + (yy - cy) * math.cos(th)
        m |= (u / a) ** 2 + (v / b) ** 2 <= 1.0
    return (m.to(torch.uint8) * 255).contiguous()


@dataclass
class SynthBatch:
    images: torch.Tensor    # [B,H,W,3] f32
    labels: torch.Tensor    # [B,H,W] i32, labels 1..L_b
    canny: torch.Tensor     # [B,H,W] u8 {0,1}
    gt: torch.Tensor        # [B,H,W] u8 {0,255}
    n_regions: np.ndarray   # [B] int64

    def to(self, device, non_blocking=False):
        return SynthBatch(self.images.to(device, non_blocking=non_blocking),
                          self.labels.to(device, non_blocking=non_blocking),
                          self.canny.to(device, non_blocking=non_blocking),
                          self.gt.to(device, non_blocking=non_blocking), self.n_regions)

    def __len__(self):
        return self.images.shape[0]

    def as_decoded(self):
        """The same batch in the form a decoder hands over: uint8 RGB (byte = round(255 * value)) and int16 labels."""
        img = self.images if self.images.dtype == torch.uint8 else (self.images * 255.0).round().clamp_(0, 255).to(torch.uint8)
        return SynthBatch(img, self.labels.to(torch.int16), self.canny, self.gt, self.n_regions)

    def as_float(self):
        """uint8 images back to the float [0,1] form ``create_region_graph`` receives (byte / 255)."""
        img = self.images.to(torch.float32) / 255.0 if self.images.dtype == torch.uint8 else self.images
        return SynthBatch(img, self.labels.to(torch.int32), self.canny, self.gt, self.n_regions)


def make_batch(first: int, count: int, seed: int = 1234, device="cpu", n_regions=None) -> SynthBatch:
    """Images ``first .. first+count-1`` of the seeded COD10K-shaped set."""
    counts = draw_region_counts(count, seed, first) if n_regions is None else np.broadcast_to(
        np.asarray(n_regions, dtype=np.int64), (count,)).copy()
    imgs, labs, cans, gts = [], [], [], []
    for k in range(count):
        i = first + k
        imgs.append(smooth_image(seed, i, device))
        labs.append(voronoi_labels(int(counts[k]), seed, i, device))
        cans.append(canny_like(seed, i, device))
        gts.append(ellipse_gt(seed, i, device))
    return SynthBatch(torch.stack(imgs), torch.stack(labs), torch.stack(cans), torch.stack(gts), counts)


def synthetic_kg(k: int, dim: int = 128, seed: int = 7, device="cpu") -> torch.Tensor:
    """``[k,dim]`` f32 rows of ``relu(randn)*0.25``: non-negative, about half zeros, like the bundled
    knowledge-graph embeddings."""
    g = torch.Generator().manual_seed(seed)
    return (torch.randn((k, dim), generator=g).clamp_min_(0.0) * 0.25).to(device)


def seeded_state_dict(shapes: dict, seed: int = 0, scale: float = 0.08) -> dict:
    """Deterministic weights independent of module construction order: tensor ``k`` is
    ``randn(shape)*scale`` from a generator seeded by ``(seed, crc32(k))``; BatchNorm running
    statistics and LayerNorm/BatchNorm scales get sane positive values."""
    import zlib
    out = {}
    for k, shp in shapes.items():
        g = torch.Generator().manual_seed((seed * 7919 + zlib.crc32(k.encode())) & 0x7FFFFFFF)
        if k.endswith("num_batches_tracked"):
            out[k] = torch.tensor(3, dtype=torch.long)
        elif k.endswith("running_var"):
            out[k] = torch.rand(shp, generator=g) * 0.5 + 0.75
        elif k.endswith("running_mean"):
            out[k] = torch.randn(shp, generator=g) * 0.1
        elif (k.startswith("bn") or ".ln_" in k or k.startswith("ln_")) and k.endswith("weight"):
            out[k] = 1.0 + torch.randn(shp, generator=g) * 0.1
        else:
            out[k] = torch.randn(shp, generator=g) * scale
    return out
