"""Oracle, track half: the 2D-3D correspondence search of the reconstruct stage.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

Reference restated here (paths relative to /root/reference):
* ``utils/sfm_reconstruction.py:14``       ``MATCHING_THRESHOLD = 2.0``
* ``utils/sfm_reconstruction.py:157-230``  ``StructureFromMotion.find_2d3d_matches``
* ``utils/sfm_reconstruction.py:211-213``  the dense test
  ``np.linalg.norm(valid_track_points[:, None] - other_img_pts, axis=2) < MATCHING_THRESHOLD``
  and ``np.where`` (row-major order of the hits)

Parity pinned: ``tests/golden/find_2d3d_vectors.npz`` holds outputs of the reference method itself,
generated in the build container by ``oracle/make_golden_tracks.py`` (matplotlib stubbed, the
object built without ``__init__``, which needs a hard-coded CSV path).
"""
from __future__ import annotations

import numpy as np

MATCHING_THRESHOLD = 2.0  # utils/sfm_reconstruction.py:14


def radius_match_numpy(track_pts, img_pts, threshold=MATCHING_THRESHOLD):
    """(track_idx, pt_idx) int64 arrays: ``np.where(dist < threshold)`` of :211-213, evaluated by
    numpy exactly as the reference does (f64, sqrt of the sum of the two squares)."""
    track_pts = np.asarray(track_pts)
    img_pts = np.asarray(img_pts)
    if len(track_pts) == 0 or len(img_pts) == 0:
        return np.zeros(0, np.int64), np.zeros(0, np.int64)
    distances = np.linalg.norm(track_pts[:, None] - img_pts, axis=2)
    t, p = np.where(distances < threshold)
    return t.astype(np.int64), p.astype(np.int64)


def radius_match_loops(track_pts, img_pts, threshold=MATCHING_THRESHOLD):
    """Scalar restatement (small cases): the same f64 operations one pair at a time."""
    ts, ps = [], []
    for i, t in enumerate(np.asarray(track_pts, dtype=np.float64)):
        for j, q in enumerate(np.asarray(img_pts)):
            dx = np.float64(t[0]) - np.float64(q[0])
            dy = np.float64(t[1]) - np.float64(q[1])
            if np.sqrt(dx * dx + dy * dy) < threshold:
                ts.append(i)
                ps.append(j)
    return np.asarray(ts, np.int64), np.asarray(ps, np.int64)


def find_2d3d_matches(point_tracks, points3D, image_id, image_pairs, load_pair, threshold=MATCHING_THRESHOLD):
    """``find_2d3d_matches`` (:157-230) with the file access factored out.

    point_tracks: list of dicts image id -> [x, y] (:150-155); points3D: list / array of 3-vectors;
    image_pairs: pair names ``pair_<id1>_<id2>`` in the order ``find_image_pairs`` returned them;
    load_pair(name) -> (pts1, pts2) or raises FileNotFoundError / ValueError (the pair is skipped,
    :221-223).  Returns (points3D [M,3], points2D [M,2]) as the reference builds them (:226-227)."""
    out3, out2 = [], []
    points3D_array = np.array(points3D)
    for pair in image_pairs:
        try:
            pts1, pts2 = load_pair(pair)
            id1, id2 = map(int, pair.split('_')[1:])
            if id1 == image_id:
                new_img_pts, other_img_pts, other_img_id = pts1, pts2, id2
            else:
                new_img_pts, other_img_pts, other_img_id = pts2, pts1, id1
            valid_idx = [k for k, tr in enumerate(point_tracks) if other_img_id in tr]
            if not valid_idx:
                continue
            valid_pts = np.array([point_tracks[k][other_img_id] for k in valid_idx])
            t, p = radius_match_numpy(valid_pts, other_img_pts, threshold)
            for ti, pi in zip(t, p):
                out3.append(points3D_array[valid_idx[ti]])
                out2.append(new_img_pts[pi])
        except (FileNotFoundError, ValueError):
            continue
    return np.array(out3), np.array(out2)
