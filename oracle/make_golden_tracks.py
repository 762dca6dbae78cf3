"""Generate tests/golden/find_2d3d_vectors.npz from the reference's own find_2d3d_matches.

Run ONCE in the build container (``python -m oracle.make_golden_tracks``), where /root/reference
exists; the fixture is committed because the reference does not travel to the GPU box.
TEST INFRASTRUCTURE ONLY.

The method under test is ``StructureFromMotion.find_2d3d_matches``
(utils/sfm_reconstruction.py:157-230).  Its module imports matplotlib / networkx through
``utils/image_selector.py``; both are stubbed when missing (they are not used on this path).  The
object is created without ``__init__`` (which reads a CSV from a hard-coded absolute path, :36) and
given exactly the attributes the method reads: matches_dir, corr_dir, constructed, point_tracks,
points3D.

Cases: synthetic scenes on a 1024 x 768 image.  Tracks are built the way the reference builds
them (:148-155, Python lists of the f32 coordinates widened to float); pair files hold f32 points,
some equal to a track point, some at distance 2 - eps / 2 / 2 + eps from one (the threshold edge),
the rest random.  One case has a missing correspondence file (skipped pair, :221-223).
"""
from __future__ import annotations

import importlib
import sys
import tempfile
import types
from pathlib import Path

import numpy as np

REF = Path("/root/reference")
OUT = Path(__file__).resolve().parent.parent / "tests" / "golden"


def load_reference_class():
    for name in ("matplotlib", "matplotlib.pyplot", "networkx"):
        try:
            importlib.import_module(name)
        except ImportError:
            sys.modules[name] = types.ModuleType(name)
    if "matplotlib" in sys.modules and not hasattr(sys.modules["matplotlib"], "pyplot"):
        sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    sys.path.insert(0, str(REF))
    try:
        mod = importlib.import_module("utils.sfm_reconstruction")
    finally:
        sys.path.remove(str(REF))
    return mod.StructureFromMotion


def edge_points(rng, base, n):
    """n f32 points around f32 points `base`: offsets of length ~0, 2 -/+ tiny, and exactly (2, 0),
    (0, 2), (1.2, 1.6) whose f64 norm of the f32 difference lands on either side of 2.0."""
    k = rng.integers(0, len(base), n)
    ang = rng.uniform(0, 2 * np.pi, n)
    r = rng.choice([0.0, 1.0, 1.9999, 2.0, 2.0001, 1.99999988, 2.00000012, 3.0], n)
    off = np.stack([r * np.cos(ang), r * np.sin(ang)], 1)
    exact = rng.integers(0, 4, n)
    off[exact == 1] = (2.0, 0.0)
    off[exact == 2] = (0.0, -2.0)
    off[exact == 3] = (1.2, 1.6)
    return (base[k].astype(np.float64) + off).astype(np.float32)


def make_case(rng, n_images, n_tracks, pts_per_pair, new_id, drop_file):
    """Reconstruction state + pair files for adding image `new_id`."""
    constructed = [i for i in range(n_images) if i != new_id and rng.random() < 0.8]
    if not constructed:
        constructed = [(new_id + 1) % n_images]
    tracks, points3D = [], []
    for _ in range(n_tracks):
        ids = rng.choice(constructed, size=min(len(constructed), int(rng.integers(1, 4))), replace=False)
        tr = {}
        for i in ids:
            pt = (rng.uniform(0, [1024, 768])).astype(np.float32)
            tr[int(i)] = pt.tolist()                       # as :151-153
        tracks.append(tr)
        points3D.append(rng.normal(size=3))
    pairs = {}
    for other in range(n_images):
        if other == new_id or rng.random() < 0.15:
            continue
        id1, id2 = (new_id, other) if rng.random() < 0.5 else (other, new_id)
        base = np.array([tr[other] for tr in tracks if other in tr], dtype=np.float32).reshape(-1, 2)
        n = int(pts_per_pair * rng.uniform(0.5, 1.5))
        if len(base):
            other_pts = np.concatenate([edge_points(rng, base, n // 2),
                                        rng.uniform(0, [1024, 768], (n - n // 2, 2)).astype(np.float32)])
        else:
            other_pts = rng.uniform(0, [1024, 768], (n, 2)).astype(np.float32)
        new_pts = rng.uniform(0, [1024, 768], (len(other_pts), 2)).astype(np.float32)
        pts1, pts2 = (new_pts, other_pts) if id1 == new_id else (other_pts, new_pts)
        pairs[f"pair_{id1}_{id2}"] = (pts1, pts2)
    dropped = None
    if drop_file and pairs:
        dropped = sorted(pairs)[len(pairs) // 2]
    return constructed, tracks, points3D, pairs, dropped


def run_reference(cls, new_id, constructed, tracks, points3D, pairs, dropped):
    with tempfile.TemporaryDirectory() as tmp:
        tmp = Path(tmp)
        (tmp / "matches").mkdir()
        (tmp / "correspondences").mkdir()
        for name, (pts1, pts2) in pairs.items():
            np.savez(tmp / "matches" / f"{name}_matches.npz", matches=np.zeros((0, 2), np.int32))
            if name != dropped:
                np.save(tmp / "correspondences" / f"{name}_pts1.npy", pts1)
                np.save(tmp / "correspondences" / f"{name}_pts2.npy", pts2)
        obj = object.__new__(cls)
        obj.matches_dir = tmp / "matches"
        obj.corr_dir = tmp / "correspondences"
        obj.constructed = [f"{i:04d}.ppm" for i in constructed]
        obj.point_tracks = tracks
        obj.points3D = points3D
        order = obj.find_image_pairs(new_id)               # directory order: part of the fixture
        obj.find_image_pairs = lambda image_id: order      # the same order inside find_2d3d_matches
        p3, p2 = obj.find_2d3d_matches(new_id)
    return order, np.asarray(p3), np.asarray(p2)


def main():
    cls = load_reference_class()
    rng = np.random.default_rng(20241104)
    out = {}
    specs = [(6, 40, 30, 2, False), (8, 300, 200, 0, True), (5, 700, 900, 4, False), (4, 3, 5, 1, False),
             (10, 1500, 400, 7, True)]
    for c, (n_images, n_tracks, ppp, new_id, drop) in enumerate(specs):
        constructed, tracks, points3D, pairs, dropped = make_case(rng, n_images, n_tracks, ppp, new_id, drop)
        order, p3, p2 = run_reference(cls, new_id, constructed, tracks, points3D, pairs, dropped)
        out[f"c{c}_new_id"] = np.int64(new_id)
        out[f"c{c}_points3D"] = np.asarray(points3D)
        # tracks as (track index, image id, x, y) rows; coordinates as the Python floats the dicts hold
        rows = [(k, i, xy[0], xy[1]) for k, tr in enumerate(tracks) for i, xy in tr.items()]
        out[f"c{c}_tracks"] = np.asarray(rows, dtype=np.float64).reshape(-1, 4)
        out[f"c{c}_n_tracks"] = np.int64(len(tracks))
        out[f"c{c}_order"] = np.asarray(order)
        out[f"c{c}_dropped"] = np.asarray(dropped or "")
        for name, (pts1, pts2) in pairs.items():
            out[f"c{c}_{name}_pts1"] = pts1
            out[f"c{c}_{name}_pts2"] = pts2
        out[f"c{c}_out3"] = p3
        out[f"c{c}_out2"] = p2
        print(f"case {c}: {len(tracks)} tracks, {len(order)} pairs, {len(p3)} 2D-3D matches, dropped={dropped}")
    out["n_cases"] = np.int64(len(specs))
    OUT.mkdir(parents=True, exist_ok=True)
    np.savez_compressed(OUT / "find_2d3d_vectors.npz", **out)
    print("wrote", OUT / "find_2d3d_vectors.npz")


if __name__ == "__main__":
    main()
