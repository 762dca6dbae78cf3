"""Generate tests/golden/* from the reference and its shipped data.

Run ONCE in the build container (``python -m oracle.make_golden``), where
/root/reference exists; the fixtures are committed because the reference does
not travel to the GPU box.  TEST INFRASTRUCTURE ONLY.

Fixtures
* bunny_fundamental.npz -- the 148 shipped ``bunny_data/fundamental/*.npz``
  (pts1, pts2, F, mask) packed, plus what the reference's OWN
  ``ImageMatcher.geometric_verification`` / ``verify_match_quality``
  (utils/find_matches.py:157-214, class loaded by file path) return for them,
  plus ``cv2.findFundamentalMat`` replayed under this container's cv2.
* knn_vectors.npz -- ``cv2.BFMatcher(NORM_L2).knnMatch(k=2)`` + the literal
  ratio loop on small seeded inputs (random, planted ties, ragged sizes).
* bunny_sift.npz -- SIFT keypoints/descriptors of the 36 bunny images by the
  recovered recipe, and the shipped match lists for the soft
  end-to-end check.
"""
from __future__ import annotations

import importlib.util
import sys
from pathlib import Path

import numpy as np

REF = Path("/root/reference")
OUT = Path(__file__).resolve().parent.parent / "tests" / "golden"


def load_reference_matcher():
    """Load utils/find_matches.py by path (``import utils`` needs matplotlib)."""
    spec = importlib.util.spec_from_file_location("ref_find_matches", REF / "utils" / "find_matches.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def pair_key(p: Path):
    _, i, j, _ = p.stem.split("_")
    return int(i), int(j)


def make_fundamental(tmp_dir: Path):
    import cv2
    ref = load_reference_matcher()
    matcher = ref.ImageMatcher(tmp_dir)      # mkdirs under tmp only
    files = sorted((REF / "bunny_data" / "fundamental").glob("pair_*_F.npz"), key=pair_key)
    names, offs = [], [0]
    P1, P2, MASK, SYM, Fs, Fcv, metrics, passed = [], [], [], [], [], [], [], []
    for f in files:
        z = np.load(f)
        pts1, pts2, F, mask = z["pts1"], z["pts2"], z["F"], z["mask"]
        res = matcher.geometric_verification(pts1, pts2, F)
        assert np.array_equal(res["inlier_mask"], mask), f.name
        m = res["metrics"]
        metrics.append([m["inliers"], m["inlier_ratio"], m["reprojection_error"],
                        m["symmetric_error"], float(m["well_distributed"])])
        passed.append(matcher.verify_match_quality(res))
        F2, _ = cv2.findFundamentalMat(pts1, pts2, cv2.FM_RANSAC, 3.0)
        names.append(f.stem[:-2])
        offs.append(offs[-1] + len(pts1))
        P1.append(pts1); P2.append(pts2); MASK.append(mask)
        SYM.append(res["symmetric_errors"]); Fs.append(F); Fcv.append(F2)
    np.savez_compressed(
        OUT / "bunny_fundamental.npz",
        names=np.array(names), offsets=np.array(offs, np.int64),
        pts1=np.concatenate(P1).astype(np.float32), pts2=np.concatenate(P2).astype(np.float32),
        mask=np.concatenate(MASK), sym=np.concatenate(SYM).astype(np.float32),
        F=np.stack(Fs), F_cv2_replay=np.stack(Fcv),
        metrics=np.array(metrics, np.float64), passed=np.array(passed),
        cv2_version=np.array(cv2.__version__))
    print("bunny_fundamental:", len(names), "pairs,", offs[-1], "points, all pass:", all(passed))


def make_knn():
    import cv2
    rng = np.random.default_rng(20240)
    cases = {}

    def run(tag, d1, d2):
        bf = cv2.BFMatcher(cv2.NORM_L2, crossCheck=False)
        ms = bf.knnMatch(d1.astype(np.float32), d2.astype(np.float32), k=2)
        idx = np.array([[m.trainIdx, n.trainIdx] for m, n in ms], np.int64)
        dist = np.array([[m.distance, n.distance] for m, n in ms], np.float32)
        good = [m for m, n in ms if m.distance < 0.75 * n.distance]
        cases[tag + "_d1"] = d1.astype(np.uint8)
        cases[tag + "_d2"] = d2.astype(np.uint8)
        cases[tag + "_idx"] = idx
        cases[tag + "_dist"] = dist
        cases[tag + "_good_q"] = np.array([m.queryIdx for m in good], np.int64)
        cases[tag + "_good_t"] = np.array([m.trainIdx for m in good], np.int64)

    run("rand", rng.integers(0, 256, (300, 128)), rng.integers(0, 256, (700, 128)))
    # low-entropy values -> many exact distance ties
    run("ties", rng.integers(0, 2, (257, 128)) * 40, rng.integers(0, 2, (513, 128)) * 40)
    d2 = rng.integers(0, 80, (400, 128))
    d2[[5, 9, 250, 399]] = d2[2]          # 5-way duplicate rows
    d1 = d2[rng.integers(0, 400, 130)] + rng.integers(0, 3, (130, 128))
    run("dups", d1, d2)
    run("tiny", rng.integers(0, 256, (1, 128)), rng.integers(0, 256, (2, 128)))
    run("ragged", rng.integers(0, 256, (52, 128)), rng.integers(0, 256, (131, 128)))
    np.savez_compressed(OUT / "knn_vectors.npz", **cases)
    print("knn_vectors:", sorted({k.split('_')[0] for k in cases}))


def make_bunny_sift():
    import cv2
    sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
    from sfm_b200 import features
    feats = features.extract_range(REF / "bunny_data", 0, 35, use_mask=True)
    ids = sorted(feats)
    offs = [0]
    for i in ids:
        offs.append(offs[-1] + len(feats[i][1]))
    out = dict(ids=np.array(ids, np.int64), offsets=np.array(offs, np.int64),
               kps=np.concatenate([feats[i][1] for i in ids]),
               desc=np.concatenate([feats[i][2] for i in ids]),
               names=np.array([feats[i][0].name for i in ids]),
               cv2_version=np.array(cv2.__version__))
    # shipped match lists (soft check) + CSV
    files = sorted((REF / "bunny_data" / "matches").glob("pair_*_matches.npz"), key=pair_key)
    mo, mq, mt, md, mn = [0], [], [], [], []
    for f in files:
        z = np.load(f)
        mn.append(f.stem[:-8])
        mq.append(z["queryIdx"]); mt.append(z["trainIdx"]); md.append(z["distance"])
        mo.append(mo[-1] + len(z["queryIdx"]))
    out.update(m_names=np.array(mn), m_offsets=np.array(mo, np.int64),
               m_query=np.concatenate(mq).astype(np.int64), m_train=np.concatenate(mt).astype(np.int64),
               m_dist=np.concatenate(md).astype(np.float64))
    np.savez_compressed(OUT / "bunny_sift.npz", **out)
    (OUT / "bunny_matching_results.csv").write_text((REF / "bunny_data" / "matching_results.csv").read_text())
    print("bunny_sift:", len(ids), "images,", offs[-1], "descriptors;", len(mn), "golden match lists")


def main():
    import tempfile
    OUT.mkdir(parents=True, exist_ok=True)
    with tempfile.TemporaryDirectory() as td:
        make_fundamental(Path(td))
    make_knn()
    make_bunny_sift()


if __name__ == "__main__":
    main()
