"""Per-image feature stage (input producer of the hot path).

Kernels start at descriptors; extraction stays on the host with OpenCV exactly
as the reference does it, but hoisted to ONCE per image instead of once per
pair (utils/find_matches.py:265-266 re-detects both images for every pair).

Recipe = the one that generated the shipped bunny artefacts:
silhouette mask as utils/find_matches.py:49-72, default ``cv2.SIFT_create()``
on the mask-ANDed grayscale image (:84-95), strict mask post-filter (:121-137).
"""
from __future__ import annotations

from pathlib import Path

import numpy as np


def find_image(image_dir: Path, idx: int):
    """The reference looks for ``%05d.jpg`` (utils/find_matches.py:334) while its
    data and its reconstruct stage use ``%04d.ppm`` (utils/sfm_reconstruction.py:574);
    accept both, reference's own naming first."""
    for name in (f"{idx:05d}.jpg", f"{idx:04d}.ppm", f"{idx:04d}.jpg", f"{idx:05d}.ppm",
                 f"{idx:04d}.png", f"{idx:05d}.png"):
        p = Path(image_dir) / name
        if p.exists():
            return p
    return None


def load_mask(silhouette_dir: Path, idx: int):
    """utils/find_matches.py:49-72 (both ``%05d`` and ``%04d`` names)."""
    import cv2
    for name in (f"{idx:05d}.pgm", f"{idx:04d}.pgm"):
        p = Path(silhouette_dir) / name
        if p.exists():
            break
    else:
        return None
    mask = cv2.imread(str(p), cv2.IMREAD_GRAYSCALE)
    if mask is None:
        return None
    _, mask = cv2.threshold(mask, 127, 255, cv2.THRESH_BINARY)
    mask = cv2.bitwise_not(mask)
    mask = cv2.morphologyEx(mask, cv2.MORPH_CLOSE, np.ones((3, 3), np.uint8))
    return mask


def detect_sift(img, mask=None):
    """(kps f32 [n,2], desc u8 [n,128]); structure of utils/find_matches.py:74-139
    with SIFT in place of FAST+ORB."""
    import cv2
    gray = cv2.cvtColor(img, cv2.COLOR_BGR2GRAY) if img.ndim == 3 else img
    if mask is not None:
        _, mask = cv2.threshold(mask, 127, 255, cv2.THRESH_BINARY)
        work = cv2.bitwise_and(gray, gray, mask=mask)
    else:
        work = gray
    kps, desc = cv2.SIFT_create().detectAndCompute(work, mask)
    if desc is None or len(kps) == 0:
        return np.zeros((0, 2), np.float32), np.zeros((0, 128), np.uint8)
    pts = np.array([k.pt for k in kps], dtype=np.float32).reshape(-1, 2)
    if mask is not None:
        x = pts[:, 0].astype(np.int64)      # int() truncation, :125
        y = pts[:, 1].astype(np.int64)
        ok = (y >= 0) & (y < mask.shape[0]) & (x >= 0) & (x < mask.shape[1])
        ok[ok] &= mask[y[ok], x[ok]] == 255
        pts, desc = pts[ok], desc[ok]
    d8 = desc.astype(np.uint8)
    assert np.array_equal(d8.astype(desc.dtype), desc), "SIFT descriptors must be integer-valued"
    return np.ascontiguousarray(pts), np.ascontiguousarray(d8)


def extract_range(data_dir, start_idx: int, end_idx: int, use_mask: bool):
    """Features for every existing image id in [start_idx, end_idx] (inclusive).

    Returns dict id -> (path, kps, desc)."""
    import cv2
    data_dir = Path(data_dir)
    out = {}
    for i in range(start_idx, end_idx + 1):
        p = find_image(data_dir / "images", i)
        if p is None:
            continue
        img = cv2.imread(str(p))
        if img is None:
            raise ValueError(f"Could not read image: {p}")       # :253
        mask = load_mask(data_dir / "silhouettes", i) if use_mask else None
        kps, desc = detect_sift(img, mask)
        out[i] = (p, kps, desc)
    return out
