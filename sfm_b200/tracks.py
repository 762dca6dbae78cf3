"""2D-3D correspondence search of the reconstruct stage on the GPU.

Mirror of ``StructureFromMotion.find_2d3d_matches`` (utils/sfm_reconstruction.py:157-230): same
arguments in spirit (the object state it reads is passed in), same outputs, same skipped-pair
behaviour.  The reference builds, for every image pair containing the new image, a dense
tracks x points distance matrix in numpy inside a Python loop (:211-213) and then appends the hits
one by one (:215-218).  Here every pair becomes one *segment* of a single ``sfm_radius_match``
call (csrc/radius.cu) and the hits are gathered with two fancy-indexing operations.

There is no CPU fallback: without the CUDA library ``Context`` raises.
"""
from __future__ import annotations

from pathlib import Path

import numpy as np

MATCHING_THRESHOLD = 2.0  # utils/sfm_reconstruction.py:14


def load_pair_from(corr_dir):
    """The reference's file access (:179-180): ``<pair>_pts1.npy`` / ``<pair>_pts2.npy``."""
    corr_dir = Path(corr_dir)

    def load(pair):
        return np.load(corr_dir / f"{pair}_pts1.npy"), np.load(corr_dir / f"{pair}_pts2.npy")
    return load


def build_segments(point_tracks, image_id, image_pairs, load_pair):
    """Host side of :176-208: per usable pair the indices of the tracks that contain the other
    image, their points there (f64), the other image's points and the new image's points."""
    by_image = {}                                        # image id -> (track indices, points), built once
    for k, tr in enumerate(point_tracks):
        for i, xy in tr.items():
            e = by_image.setdefault(i, ([], []))
            e[0].append(k)
            e[1].append(xy)
    segs = []
    for pair in image_pairs:
        try:
            pts1, pts2 = load_pair(pair)
            id1, id2 = map(int, pair.split('_')[1:])
            if id1 == image_id:
                new_pts, other_pts, other_id = pts1, pts2, id2
            else:
                new_pts, other_pts, other_id = pts2, pts1, id1
            if other_id not in by_image:
                continue                                 # :203-204
            idx, xy = by_image[other_id]
            tp = np.array(xy)                            # :207, float64 from the Python floats
            other_pts = np.asarray(other_pts)
            if tp.ndim != 2 or tp.shape[1] != 2 or (other_pts.size and (other_pts.ndim != 2 or other_pts.shape[1] != 2)):
                raise ValueError("points must be [n, 2]")    # numpy's broadcast error in the reference
            segs.append((np.asarray(idx, np.int64), tp.astype(np.float64, copy=False),
                         other_pts.reshape(-1, 2), np.asarray(new_pts)))
        except (FileNotFoundError, ValueError):          # :221-223
            continue
    return segs


def find_2d3d_matches(ctx, point_tracks, points3D, image_id, image_pairs, load_pair,
                      threshold=MATCHING_THRESHOLD):
    """(points3D [M,3], points2D [M,2]) exactly as :226-230 returns them.

    ctx: ``sfm_b200.lib.Context``; image_pairs: pair names in ``find_image_pairs`` order (:551-579);
    load_pair: name -> (pts1, pts2), e.g. ``load_pair_from(corr_dir)``."""
    torch = ctx.torch
    segs = build_segments(point_tracks, image_id, image_pairs, load_pair)
    if not segs:
        return np.array([]), np.array([])
    track_off = np.zeros(len(segs) + 1, np.int64)
    pt_off = np.zeros(len(segs) + 1, np.int64)
    for s, (idx, tp, other, _) in enumerate(segs):
        track_off[s + 1] = track_off[s] + len(tp)
        pt_off[s + 1] = pt_off[s] + len(other)
    tracks = np.concatenate([sg[1] for sg in segs])
    pts = np.concatenate([sg[2].astype(np.float64) for sg in segs])      # f32 -> f64 is exact
    d_tracks = torch.from_numpy(np.ascontiguousarray(tracks)).to(ctx.device)
    d_pts = torch.from_numpy(np.ascontiguousarray(pts)).to(ctx.device)
    seg_off, t_idx, p_idx = ctx.radius_match(track_off, d_tracks, pt_off, d_pts, threshold)
    if int(seg_off[-1]) == 0:
        return np.array([]), np.array([])
    t_idx = t_idx.cpu().numpy()
    p_idx = p_idx.cpu().numpy()
    points3D_array = np.array(points3D)
    out3, out2 = [], []
    for s, (idx, _, _, new_pts) in enumerate(segs):
        a, b = int(seg_off[s]), int(seg_off[s + 1])
        if a == b:
            continue
        out3.append(points3D_array[idx[t_idx[a:b]]])
        out2.append(new_pts[p_idx[a:b]])
    return np.concatenate(out3), np.concatenate(out2)
