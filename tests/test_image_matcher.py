"""Host-side mirror of the reference's ImageMatcher (utils/find_matches.py): interface and
error behaviour on CPU, end-to-end parity with the oracle on the GPU."""
import inspect

import numpy as np
import pytest

from sfm_b200 import find_matches as fmx
from sfm_b200 import lib as L


def test_interface_mirrors_reference_methods(tmp_path):
    m = fmx.ImageMatcher(tmp_path)
    for d in ("matches", "fundamental", "correspondences", "visualizations"):       # :21-22
        assert (tmp_path / d).is_dir()
    assert m.results == []
    sig = {
        "process_image_range": ["start_idx", "end_idx", "mask"],                     # :329
        "process_image_pair": ["img1_path", "img2_path", "pair_name", "mask"],        # :246
        "match_features": ["desc1", "desc2"],                                         # :141
        "geometric_verification": ["pts1", "pts2", "F", "threshold"],                 # :157
        "verify_match_quality": ["geometric_results", "min_inliers", "min_ratio", "max_error"],   # :203
        "save_pair_data": ["pair_name", "pts1", "pts2", "F", "inlier_mask", "matches"],            # :312
        "save_results": ["output_csv"],                                               # :352
        "load_mask": ["idx"], "detect_features": ["img", "mask"],
    }
    for name, args in sig.items():
        assert list(inspect.signature(getattr(m, name)).parameters) == args
    p = inspect.signature(m.geometric_verification).parameters
    assert p["threshold"].default == 3.0
    q = inspect.signature(m.verify_match_quality).parameters
    assert (q["min_inliers"].default, q["min_ratio"].default, q["max_error"].default) == (15, 0.3, 2.0)


def test_quality_gate_and_empty_results(tmp_path):
    m = fmx.ImageMatcher(tmp_path)
    good = {"metrics": {"inliers": 15, "inlier_ratio": 0.3, "reprojection_error": 2.0, "well_distributed": True}}
    assert m.verify_match_quality(good)
    for k, v in (("inliers", 14), ("inlier_ratio", 0.29), ("reprojection_error", 2.01), ("well_distributed", False)):
        bad = {"metrics": dict(good["metrics"], **{k: v})}
        assert not m.verify_match_quality(bad)
    m.process_image_range(0, 5)                       # no images: nothing to do, no error (:334-337)
    assert m.results == []
    with pytest.raises(KeyError):                     # reference behaviour with zero rows (:360)
        m.save_results(tmp_path / "matching_results.csv")


def test_no_cpu_fallback(tmp_path):
    torch = pytest.importorskip("torch")
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    m = fmx.ImageMatcher(tmp_path)
    d = np.zeros((4, 128), np.uint8)
    with pytest.raises(L.SfmError):
        m.match_features(d, d)
    feats = {0: (tmp_path / "0000.ppm", np.zeros((4, 2), np.float32), d),
             1: (tmp_path / "0001.ppm", np.zeros((4, 2), np.float32), d)}
    with pytest.raises(L.SfmError):
        m.process_features(feats, 0, 1)


@pytest.mark.gpu
def test_bunny_range_files_and_parity_with_oracle(tmp_path, golden_dir):
    """The 36 bunny images' real SIFT features through process_features: output files have the
    reference's schema (:312-327), match lists equal the oracle's exactly, and the pass/skip
    decision follows the oracle except at the RANSAC gate boundary."""
    pytest.importorskip("torch")
    import pandas as pd
    from oracle import fundamental as ofm
    from oracle import knn as oknn
    z = np.load(golden_dir / "bunny_sift.npz")
    off, ids = z["offsets"], z["ids"]
    feats = {int(i): (tmp_path / "images" / f"{int(i):04d}.ppm", z["kps"][off[n]:off[n + 1]],
                      z["desc"][off[n]:off[n + 1]]) for n, i in enumerate(ids)}
    m = fmx.ImageMatcher(tmp_path)
    lo, hi = int(ids.min()), int(ids.max())
    m.process_features(feats, lo, hi)
    res, names = m.last_batch, m.last_names
    assert len(names) == 36 * 35 // 2
    n_ok = int((res.status == L.PAIR_OK).sum())
    assert len(m.results) == n_ok and n_ok > 100
    m.save_results(tmp_path / "matching_results.csv")
    df = pd.read_csv(tmp_path / "matching_results.csv")
    assert list(df.columns) == fmx.CSV_COLUMNS
    assert df["img1"].iloc[0].endswith(".ppm")
    by_name = {r["pair_name"]: r for r in m.results}
    for p, (_, _, name) in enumerate(names):
        i, j = (int(x) for x in name.split("_")[1:])
        q, t, d = oknn.match_features(feats[i][2], feats[j][2])
        if res.status[p] == L.PAIR_TOO_FEW_MATCHES:
            assert len(q) < 5
            continue
        assert res.count[p] == len(q)
        if res.status[p] != L.PAIR_OK:
            assert not (tmp_path / "matches" / f"{name}_matches.npz").exists()      # only passing pairs are saved
            continue
        mz = np.load(tmp_path / "matches" / f"{name}_matches.npz")
        assert mz["queryIdx"].dtype == np.int64 and mz["trainIdx"].dtype == np.int64
        assert mz["distance"].dtype == np.float64 and mz["inlier_mask"].dtype == bool
        assert np.array_equal(mz["queryIdx"], q) and np.array_equal(mz["trainIdx"], t)
        assert np.array_equal(mz["distance"], d.astype(np.float64))
        fz = np.load(tmp_path / "fundamental" / f"{name}_F.npz")
        assert fz["F"].shape == (3, 3) and fz["F"].dtype == np.float64
        assert fz["pts1"].dtype == np.float32 and fz["pts1"].shape == (len(q), 2)
        assert np.array_equal(fz["pts1"], feats[i][1][q]) and np.array_equal(fz["pts2"], feats[j][1][t])
        v = ofm.geometric_verification(fz["pts1"], fz["pts2"], fz["F"])              # oracle on OUR F
        assert np.array_equal(v["inlier_mask"], fz["mask"]) and ofm.verify_match_quality(v)
        c1 = np.load(tmp_path / "correspondences" / f"{name}_pts1.npy")
        assert np.array_equal(c1, fz["pts1"][fz["mask"]])
        row = by_name[name]
        assert row["num_matches"] == len(q) and row["num_inliers"] == int(fz["mask"].sum())
        assert abs(row["inlier_ratio"] - v["metrics"]["inlier_ratio"]) < 1e-12
        assert np.isclose(row["reprojection_error"], v["metrics"]["reprojection_error"], rtol=1e-4)


@pytest.mark.gpu
def test_match_pairs_chunked_equals_single_chunk():
    """PairEngine.match_pairs with a tiny chunk size (many chunks, results fetched behind the next
    chunk's compute into one pinned arena that has to grow) returns exactly the single-chunk result."""
    from sfm_b200 import synth
    from sfm_b200.engine import PairEngine
    sizes = [900, 400, 1300, 700, 650, 1000, 300, 820]
    s = synth.make_image_set(len(sizes), sizes, seed=21, shared_frac=0.6, tie_frac=0.02)
    pairs = np.array([(i, j) for i in range(len(sizes)) for j in range(i + 1, len(sizes))], np.int32)
    # later pairs carry more matches than the first ones would predict: forces the arena to grow
    order = np.argsort([sizes[i] for i, _ in pairs], kind="stable")
    pairs = pairs[order]
    one = PairEngine(0)
    one.upload_images(s.desc, s.kps, s.offsets.numpy())
    a = one.match_pairs(pairs, only_ok=True)       # passing pairs only: every packed array is defined
    many = PairEngine(0, max_rows_per_chunk=1500)
    many.upload_images(s.desc, s.kps, s.offsets.numpy())
    assert sum(1 for _ in many._chunks(pairs)) > 8
    b = many.match_pairs(pairs, only_ok=True)
    assert int(a.pack_off[-1]) > 100 and int((a.status == 0).sum()) >= 10
    for f in ("status", "count", "F", "pack_off", "query_idx", "train_idx", "sq_dist", "inlier_mask", "sym_err",
              "pts1", "pts2"):
        x, y = getattr(a, f), getattr(b, f)
        assert x.dtype == y.dtype and x.shape == y.shape and np.array_equal(x, y, equal_nan=True), f
    assert a.metrics.tobytes() == b.metrics.tobytes()
