"""Deterministic synthetic SIFT-like image sets.

A 3-D point cloud in the bunny bounding box is viewed by cameras on a ring with
the reference's intrinsics (utils/sfm_reconstruction.py:40-49), so image pairs
have true two-view geometry and RANSAC has real inliers.  Every scene point has
a base descriptor with SIFT's value statistics (integer 0..255, L2 norm ~512,
~25 % zeros, heavy tail); each view adds integer noise.  Rows that are not
shared scene points are image-unique clutter.

Only tensor plumbing: runs on CPU or CUDA torch; results are u8 / f32 tensors.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch

BASE_SEED = 20240
K_INTR = (1228.0, 1228.0, 512.0, 384.0)   # fx, fy, cx, cy
IMG_W, IMG_H = 1024, 768


def sift_like(n: int, gen: torch.Generator, device) -> torch.Tensor:
    """float32 [n,128] un-noised base descriptors with SIFT statistics (x512 scale)."""
    g = torch.randn(n, 128, generator=gen, device=device)
    u = torch.rand(n, 128, generator=gen, device=device)
    mag = torch.exp(1.6 * g) * (u > 0.22)
    mag = mag / mag.norm(dim=1, keepdim=True).clamp_min(1e-6)
    mag = mag.clamp_max(0.2)
    mag = mag / mag.norm(dim=1, keepdim=True).clamp_min(1e-6)
    return mag * 512.0


def quantise(x: torch.Tensor) -> torch.Tensor:
    return x.round().clamp_(0, 255).to(torch.uint8)


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
    c = torch.tensor([radius * math.cos(theta), height, radius * math.sin(theta)])
    target = torch.tensor([0.0, 0.0, 3.75])
    z = target - c
    z = z / z.norm()
    up = torch.tensor([0.0, 1.0, 0.0])
    x = torch.linalg.cross(up, z)
    x = x / x.norm()
    y = torch.linalg.cross(z, x)
    R = torch.stack([x, y, z])
    return R, -R @ c


def make_image_set(n_images: int, n_desc, seed: int = BASE_SEED, device="cpu",
                   shared_frac: float = 0.5, view_width: float = 1.2,
                   noise_sigma: float = 6.0, tie_frac: float = 0.0,
                   arc: float = 2 * math.pi) -> ImageSet:
    """n_desc: int (uniform) or a list of per-image counts (ragged)."""
    device = torch.device(device)
    counts = [int(n_desc)] * n_images if isinstance(n_desc, int) else [int(c) for c in n_desc]
    assert len(counts) == n_images
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    nmax = max(max(counts), 8)
    n_scene = max(4 * nmax, 64)
    # scene points in the bunny bounding box (bunny_data/README.txt:24-28)
    lo = torch.tensor([-7.5, -10.0, -7.5], device=device)
    hi = torch.tensor([7.5, 10.0, 15.0], device=device)
    X = lo + (hi - lo) * torch.rand(n_scene, 3, generator=gen, device=device)
    phi = arc * torch.rand(n_scene, generator=gen, device=device)   # visibility centre
    base = sift_like(n_scene, gen, device)
    fx, fy, cx, cy = K_INTR
    descs, kps = [], []
    for i, n in enumerate(counts):
        theta = arc * i / max(n_images, 1)
        R, t = _ring_camera(theta)
        R, t = R.to(device), t.to(device)
        dphi = torch.remainder(phi - theta + math.pi, 2 * math.pi) - math.pi
        vis = torch.nonzero(dphi.abs() < view_width).flatten()
        n_sh = min(int(shared_frac * n), vis.numel())
        perm = torch.randperm(vis.numel(), generator=gen, device=device)[:n_sh]
        sel = vis[perm]
        Xc = X[sel] @ R.T + t
        uv = torch.stack([fx * Xc[:, 0] / Xc[:, 2] + cx, fy * Xc[:, 1] / Xc[:, 2] + cy], 1)
        uv = uv + 0.5 * torch.randn(n_sh, 2, generator=gen, device=device)
        d_sh = base[sel] + noise_sigma * torch.randn(n_sh, 128, generator=gen, device=device)
        n_cl = n - n_sh
        d_cl = sift_like(n_cl, gen, device) if n_cl > 0 else base[:0]
        uv_cl = torch.rand(n_cl, 2, generator=gen, device=device) * torch.tensor(
            [float(IMG_W), float(IMG_H)], device=device)
        d = quantise(torch.cat([d_sh, d_cl]))
        k = torch.cat([uv, uv_cl]).float()
        if n > 0:
            order = torch.randperm(n, generator=gen, device=device)
            d, k = d[order], k[order]
            n_tie = int(tie_frac * n)
            if n_tie > 0 and n > 2 * n_tie:
                # duplicated rows exercise the lowest-index tie rule
                src = torch.randint(0, n, (n_tie,), generator=gen, device=device)
                dst = torch.randint(0, n, (n_tie,), generator=gen, device=device)
                d[dst] = d[src]
        descs.append(d)
        kps.append(k)
    offs = torch.zeros(n_images + 1, dtype=torch.int64)
    offs[1:] = torch.tensor(counts, dtype=torch.int64).cumsum(0)
    return ImageSet(torch.cat(descs).contiguous(), torch.cat(kps).contiguous(), offs)


def all_pairs(n_images: int) -> torch.Tensor:
    """i32 [P,2], i-major, j ascending: the loop order of
    utils/find_matches.py:332-341."""
    iu = torch.triu_indices(n_images, n_images, offset=1)
    return iu.T.contiguous().to(torch.int32)
