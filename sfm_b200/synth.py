"""Deterministic synthetic SIFT-like image sets.

A 3-D point cloud in the bunny bounding box is viewed by cameras on a ring with
the reference's intrinsics (utils/sfm_reconstruction.py:40-49), so image pairs
have true two-view geometry and RANSAC has real inliers.  Every scene point has
a base descriptor with SIFT's value statistics (integer 0..255, L2 norm ~512,
~25 % zeros, heavy tail); each view adds integer noise.  Rows that are not
shared scene points are image-unique clutter.

Everything is drawn on the HOST with ``numpy.random.default_rng``: the
scene from ``(seed,)``, image i from ``(seed, 1 + i)``.  An image is therefore a pure
function of (seed, set parameters, i): the CPU arm, the CUDA arm and every rank see
byte-identical inputs, and any subset of images can be generated on its own
(``make_images``).  ``device`` only says where the finished arrays are placed.
"""
from __future__ import annotations

import math
import os
from concurrent.futures import ThreadPoolExecutor
from dataclasses import dataclass

import numpy as np
import torch

BASE_SEED = 20240
K_INTR = (1228.0, 1228.0, 512.0, 384.0)   # fx, fy, cx, cy
IMG_W, IMG_H = 1024, 768
BOX_LO = np.array([-7.5, -10.0, -7.5], np.float32)     # bunny_data/README.txt:24-28
BOX_HI = np.array([7.5, 10.0, 15.0], np.float32)


def sift_like(n: int, rng: np.random.Generator) -> np.ndarray:
    """float32 [n,128] un-noised base descriptors with SIFT statistics (x512 scale)."""
    g = rng.standard_normal((n, 128), dtype=np.float32)
    u = rng.random((n, 128), dtype=np.float32)
    mag = np.exp(np.float32(1.6) * g) * (u > np.float32(0.22))
    mag /= np.maximum(np.linalg.norm(mag, axis=1, keepdims=True), np.float32(1e-6))
    np.minimum(mag, np.float32(0.2), out=mag)
    mag /= np.maximum(np.linalg.norm(mag, axis=1, keepdims=True), np.float32(1e-6))
    mag *= np.float32(512.0)
    return mag


def quantise(x: np.ndarray) -> np.ndarray:
    return np.clip(np.rint(x), 0, 255).astype(np.uint8)


@dataclass
class ImageSet:
    desc: torch.Tensor      # u8 [sum n_i, 128] packed row-major
    kps: torch.Tensor       # f32 [sum n_i, 2]
    offsets: torch.Tensor   # i64 [N+1] (CPU)

    @property
    def n_images(self) -> int:
        return self.offsets.numel() - 1

    def image(self, i: int):
        a, b = int(self.offsets[i]), int(self.offsets[i + 1])
        return self.desc[a:b], self.kps[a:b]


def _ring_camera(theta: float, radius: float = 45.0, height: float = 8.0):
    """World->camera rotation R (3x3) and translation t for a camera on a ring
    looking at the centre of the bunny box."""
    c = np.array([radius * math.cos(theta), height, radius * math.sin(theta)], np.float64)
    target = np.array([0.0, 0.0, 3.75])
    z = target - c
    z /= np.linalg.norm(z)
    up = np.array([0.0, 1.0, 0.0])
    x = np.cross(up, z)
    x /= np.linalg.norm(x)
    y = np.cross(z, x)
    R = np.stack([x, y, z])
    return R.astype(np.float32), (-R @ c).astype(np.float32)


@dataclass
class Scene:
    """What all images of a set share: 3-D points, their visibility centres and base descriptors."""
    seed: int
    n_images: int
    counts: list
    X: np.ndarray
    phi: np.ndarray
    base: np.ndarray
    shared_frac: float
    view_width: float
    noise_sigma: float
    tie_frac: float
    arc: float


def make_scene(n_images: int, n_desc, seed: int = BASE_SEED, shared_frac: float = 0.5, view_width: float = 1.2,
               noise_sigma: float = 6.0, tie_frac: float = 0.0, arc: float = 2 * math.pi) -> Scene:
    counts = [int(n_desc)] * n_images if isinstance(n_desc, (int, np.integer)) else [int(c) for c in n_desc]
    assert len(counts) == n_images
    rng = np.random.default_rng((int(seed),))
    nmax = max(max(counts, default=0), 8)
    n_scene = max(4 * nmax, 64)
    X = BOX_LO + (BOX_HI - BOX_LO) * rng.random((n_scene, 3), dtype=np.float32)
    phi = np.float32(arc) * rng.random(n_scene, dtype=np.float32)     # visibility centre
    base = sift_like(n_scene, rng)
    return Scene(int(seed), n_images, counts, X, phi, base, shared_frac, view_width, noise_sigma, tie_frac, arc)


def make_image(sc: Scene, i: int):
    """(desc u8 [n,128], kps f32 [n,2]) of image i: a pure function of the scene and i."""
    n = sc.counts[i]
    rng = np.random.default_rng((sc.seed, 1 + int(i)))
    theta = sc.arc * i / max(sc.n_images, 1)
    R, t = _ring_camera(theta)
    dphi = np.remainder(sc.phi - np.float32(theta) + np.float32(math.pi), np.float32(2 * math.pi)) - np.float32(math.pi)
    vis = np.flatnonzero(np.abs(dphi) < sc.view_width)
    n_sh = min(int(sc.shared_frac * n), len(vis))
    sel = vis[rng.permutation(len(vis))[:n_sh]]
    Xc = sc.X[sel] @ R.T + t
    fx, fy, cx, cy = K_INTR
    uv = np.stack([fx * Xc[:, 0] / Xc[:, 2] + cx, fy * Xc[:, 1] / Xc[:, 2] + cy], 1).astype(np.float32)
    uv += np.float32(0.5) * rng.standard_normal((n_sh, 2), dtype=np.float32)
    d_sh = sc.base[sel] + np.float32(sc.noise_sigma) * rng.standard_normal((n_sh, 128), dtype=np.float32)
    n_cl = n - n_sh
    d_cl = sift_like(n_cl, rng) if n_cl > 0 else np.zeros((0, 128), np.float32)
    uv_cl = rng.random((n_cl, 2), dtype=np.float32) * np.array([IMG_W, IMG_H], np.float32)
    d = quantise(np.concatenate([d_sh, d_cl]))
    k = np.concatenate([uv, uv_cl]).astype(np.float32)
    if n > 0:
        order = rng.permutation(n)
        d, k = d[order], k[order]
        n_tie = int(sc.tie_frac * n)
        if n_tie > 0 and n > 2 * n_tie:
            # duplicated rows exercise the lowest-index tie rule
            src = rng.integers(0, n, n_tie)
            dst = rng.integers(0, n, n_tie)
            d[dst] = d[src]
    return np.ascontiguousarray(d.reshape(n, 128)), np.ascontiguousarray(k.reshape(n, 2))


def make_images(sc: Scene, ids, threads: int | None = None):
    """[(desc, kps)] for the given image ids (thread pool: numpy releases the GIL in its kernels)."""
    ids = list(ids)
    threads = threads or min(16, os.cpu_count() or 1)
    if len(ids) <= 2 or threads <= 1:
        return [make_image(sc, i) for i in ids]
    with ThreadPoolExecutor(max_workers=threads) as pool:
        return list(pool.map(lambda i: make_image(sc, i), ids))


def make_image_set(n_images: int, n_desc, seed: int = BASE_SEED, device="cpu",
                   shared_frac: float = 0.5, view_width: float = 1.2,
                   noise_sigma: float = 6.0, tie_frac: float = 0.0,
                   arc: float = 2 * math.pi, pin: bool = False) -> ImageSet:
    """n_desc: int (uniform) or a list of per-image counts (ragged).  The arrays are generated on the
    host and placed on `device` (identical bytes wherever they end up)."""
    sc = make_scene(n_images, n_desc, seed, shared_frac, view_width, noise_sigma, tie_frac, arc)
    imgs = make_images(sc, range(n_images))
    offs = torch.zeros(n_images + 1, dtype=torch.int64)
    offs[1:] = torch.tensor(sc.counts, dtype=torch.int64).cumsum(0)
    total = int(offs[-1])
    desc = torch.empty((total, 128), dtype=torch.uint8, pin_memory=pin)
    kps = torch.empty((total, 2), dtype=torch.float32, pin_memory=pin)
    dn, kn = desc.numpy(), kps.numpy()
    for i, (d, k) in enumerate(imgs):
        a, b = int(offs[i]), int(offs[i + 1])
        dn[a:b] = d
        kn[a:b] = k
    device = torch.device(device)
    if device.type != "cpu":
        desc, kps = desc.to(device), kps.to(device)
    return ImageSet(desc.contiguous(), kps.contiguous(), offs)


# Named sets of SURVEY.md 8d beside the uniform BASELINE shapes.
RAGGED_COUNTS = [1, 2, 7, 14, 15, 52, 97, 128, 200, 257, 333, 400, 500]


def make_ragged_set(seed: int = BASE_SEED + 6, device="cpu", tie_frac: float = 0.02) -> ImageSet:
    """The ragged set: counts from {1, 2, 7, 14, 15} (padding and skip paths) and {52..500} (bunny range)."""
    return make_image_set(len(RAGGED_COUNTS), RAGGED_COUNTS, seed=seed, device=device, tie_frac=tie_frac,
                          arc=1.5)


def make_tie_set(n_images: int = 4, n_desc: int = 1024, seed: int = BASE_SEED + 7, device="cpu") -> ImageSet:
    """The tie set: 5 % duplicated rows per image (>= 1 % asked by SURVEY 8d) for the lowest-index rule."""
    return make_image_set(n_images, n_desc, seed=seed, device=device, tie_frac=0.05, arc=1.0)


def all_pairs(n_images: int) -> torch.Tensor:
    """i32 [P,2], i-major, j ascending: the loop order of
    utils/find_matches.py:332-341."""
    iu = torch.triu_indices(n_images, n_images, offset=1)
    return iu.T.contiguous().to(torch.int32)
