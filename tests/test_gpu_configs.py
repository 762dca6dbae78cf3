"""GPU parity on the configurations BASELINE.json names, through the public classes (PairEngine /
ImageMatcher) and the C ABI under them:

* mutual-NN end to end (README.md:108; oracle.knn.match_features(mutual=True) == BFMatcher crossCheck
  on top of the ratio test) on the 630 bunny pairs and on an 8k-descriptor pair;
* the reference's own set of passing pairs (bunny_data/matching_results.csv, 148 pairs) against ours:
  the symmetric difference is printed and every differing pair must sit on a gate boundary;
* the whole 36 x 8192 bench set (BASELINE configs[1], the bit-exactness gate): all 630 pairs' kNN + ratio
  lists against live cv2.BFMatcher.knnMatch (utils/find_matches.py:144-153);
* RANSAC under the SURVEY 8c envelope on 210 synthetic 8k-config pairs against live cv2.findFundamentalMat
  (utils/find_matches.py:282) -- bunny has at most 235 matches per pair, the bench workload a median of 450;
* an image set without a single feature (reference: every pair is skipped at :268, nothing raises).
"""
import csv

import numpy as np
import pytest

from oracle import fundamental as ofm
from oracle import knn as oknn

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _bunny(golden_dir):
    z = np.load(golden_dir / "bunny_sift.npz")
    off = z["offsets"]
    descs = [z["desc"][off[i]:off[i + 1]] for i in range(len(off) - 1)]
    kps = [z["kps"][off[i]:off[i + 1]] for i in range(len(off) - 1)]
    return z, off, descs, kps


@pytest.mark.parametrize("mutual", [True, "full"])
def test_mutual_engine_bunny_all_pairs(golden_dir, mutual):
    """mutual=True: the sparse form (sfm_mutual_filter: only the surviving matches' train descriptors go
    back against the query image); "full": a second exhaustive kNN pass over the reversed pairs."""
    from sfm_b200 import lib as L, synth
    from sfm_b200.engine import PairEngine
    z, off, descs, kps = _bunny(golden_dir)
    eng = PairEngine(0, mutual=mutual)
    eng.upload_images(torch.from_numpy(z["desc"]), torch.from_numpy(z["kps"].astype(np.float32)), off.astype(np.int64))
    pairs = synth.all_pairs(len(descs)).numpy()
    res = eng.match_pairs(pairs, only_ok=False)
    fewer = 0
    for p, (i, j) in enumerate(pairs):
        q, t, d = oknn.match_features(descs[i], descs[j], mutual=True)
        q0, _, _ = oknn.match_features(descs[i], descs[j])
        fewer += len(q) < len(q0)
        sl = res.pair_slice(p)
        assert res.count[p] == len(q), (i, j)
        assert np.array_equal(res.query_idx[sl], q) and np.array_equal(res.train_idx[sl], t), (i, j)
        assert np.array_equal(res.distance(p), d)
        assert np.array_equal(res.pts1[sl], kps[i][q]) and np.array_equal(res.pts2[sl], kps[j][t])
        if res.status[p] in (L.PAIR_OK, L.PAIR_QUALITY_FAIL):
            v = ofm.geometric_verification(res.pts1[sl], res.pts2[sl], res.F[p])
            assert np.array_equal(v["inlier_mask"], res.inlier_mask[sl])
            assert ofm.verify_match_quality(v) == (res.status[p] == L.PAIR_OK)
    assert fewer > 50          # the mutual check really removes matches on this data


def test_mutual_image_matcher_writes_mutual_lists(tmp_path, golden_dir):
    from sfm_b200 import find_matches as fmx
    from sfm_b200 import lib as L
    z, off, descs, kps = _bunny(golden_dir)
    ids = z["ids"]
    feats = {int(i): (tmp_path / "images" / f"{int(i):04d}.ppm", kps[n], descs[n]) for n, i in enumerate(ids)}
    m = fmx.ImageMatcher(tmp_path, mutual=True)
    m.process_features(feats, int(ids.min()), int(ids.max()))
    res, names = m.last_batch, m.last_names
    n_ok = 0
    for p, (_, _, name) in enumerate(names):
        if res.status[p] != L.PAIR_OK:
            continue
        n_ok += 1
        i, j = (int(x) for x in name.split("_")[1:])
        q, t, d = oknn.match_features(feats[i][2], feats[j][2], mutual=True)
        mz = np.load(tmp_path / "matches" / f"{name}_matches.npz")
        assert np.array_equal(mz["queryIdx"], q) and np.array_equal(mz["trainIdx"], t)
        assert np.array_equal(mz["distance"], d.astype(np.float64))
    assert n_ok > 80 and len(m.results) == n_ok
    # the same pair through the one-pair entry point of the reference interface
    i, j = (int(x) for x in names[0][2].split("_")[1:])
    got = m.match_features(feats[i][2], feats[j][2])
    q, t, _ = oknn.match_features(feats[i][2], feats[j][2], mutual=True)
    assert [g.queryIdx for g in got] == q.tolist() and [g.trainIdx for g in got] == t.tolist()


@pytest.mark.parametrize("mutual", [True, "full"])
def test_mutual_8k_pair(mutual):
    from sfm_b200 import synth
    from sfm_b200.engine import PairEngine
    s = synth.make_image_set(3, 8192, seed=synth.BASE_SEED + 9, tie_frac=0.01)
    eng = PairEngine(0, mutual=mutual)
    eng.upload_images(s.desc, s.kps, s.offsets.numpy())
    pairs = np.array([(0, 1), (2, 1)], np.int32)
    res = eng.match_pairs(pairs, only_ok=False)
    for p, (i, j) in enumerate(pairs):
        d1, d2 = s.image(int(i))[0].numpy(), s.image(int(j))[0].numpy()
        q, t, d = oknn.match_features(d1, d2, mutual=True)
        sl = res.pair_slice(p)
        assert len(q) > 100
        assert np.array_equal(res.query_idx[sl], q) and np.array_equal(res.train_idx[sl], t)
        assert np.array_equal(res.distance(p), d)


def test_mutual_sparse_many_chunks_and_duplicates():
    """The sparse mutual check over several chunks, on a set with 5 % duplicated rows (the reverse side's
    lowest-index rule decides) and images of very different sizes."""
    from sfm_b200 import synth
    from sfm_b200.engine import PairEngine
    sizes = [900, 400, 1300, 700, 257, 1000]
    s = synth.make_image_set(len(sizes), sizes, seed=23, shared_frac=0.6, tie_frac=0.05)
    pairs = np.array([(i, j) for i in range(len(sizes)) for j in range(len(sizes)) if i != j], np.int32)
    eng = PairEngine(0, mutual=True, max_rows_per_chunk=2500)
    eng.upload_images(s.desc, s.kps, s.offsets.numpy())
    assert sum(1 for _ in eng._chunks(pairs)) > 4
    res = eng.match_pairs(pairs, only_ok=False)
    for p, (i, j) in enumerate(pairs):
        q, t, d = oknn.match_features(s.image(int(i))[0].numpy(), s.image(int(j))[0].numpy(), mutual=True)
        sl = res.pair_slice(p)
        assert np.array_equal(res.query_idx[sl], q) and np.array_equal(res.train_idx[sl], t), (i, j)
        assert np.array_equal(res.distance(p), d)


def test_passing_pair_set_vs_reference_csv(golden_dir, capsys):
    """bunny_data/matching_results.csv (the reference's own run) lists 148 passing pairs.  Ours may differ
    only on pairs where a quality gate (utils/find_matches.py:203-214) is marginal for the side that fails
    it: inliers within 3 of the gate of 15, point spread within 5 px of the 20 px gate (the spread is a
    standard deviation over the inlier set, so it moves with every inlier gained or lost), error within 0.3
    of 2 px, ratio within 0.05 of 0.3.  The reference side of a pair it does not list is re-run here with
    its own calls (cv2.findFundamentalMat + verification, oracle/fundamental.py) on the same matches."""
    from sfm_b200 import lib as L, synth
    from sfm_b200.engine import PairEngine
    z, off, descs, kps = _bunny(golden_dir)
    ids = [int(i) for i in z["ids"]]
    eng = PairEngine(0)
    eng.upload_images(torch.from_numpy(z["desc"]), torch.from_numpy(z["kps"].astype(np.float32)), off.astype(np.int64))
    pairs = synth.all_pairs(len(descs)).numpy()
    res = eng.match_pairs(pairs, only_ok=False)
    name_of = {p: f"pair_{ids[i]}_{ids[j]}" for p, (i, j) in enumerate(pairs)}
    ours = {name_of[p] for p in range(len(pairs)) if res.status[p] == L.PAIR_OK}
    with open(golden_dir / "bunny_matching_results.csv") as f:
        ref = {r["pair_name"]: r for r in csv.DictReader(f)}
    assert len(ref) == 148
    only_ours, only_ref = sorted(ours - set(ref)), sorted(set(ref) - ours)
    by_name = {v: k for k, v in name_of.items()}

    def gates(v, p1, p2):
        m = v["metrics"]
        inl = v["inlier_mask"]
        spread = float(min(p1[inl].std(0).min(), p2[inl].std(0).min())) if inl.sum() > 1 else 0.0
        return dict(inliers=int(m["inliers"]), ratio=float(m["inlier_ratio"]), err=float(m["reprojection_error"]),
                    spread=spread)

    def failing_gate_is_marginal(g):
        fails = []
        if g["inliers"] < 15:
            fails.append(g["inliers"] >= 12)
        if g["ratio"] < 0.3:
            fails.append(g["ratio"] >= 0.25)
        if g["err"] > 2.0:
            fails.append(g["err"] <= 2.3)
        if g["spread"] <= 20.0:
            fails.append(g["spread"] >= 15.0)
        return bool(fails) and all(fails)

    report, bad = [], []
    for name in only_ours + only_ref:
        p = by_name[name]
        sl = res.pair_slice(p)
        p1, p2 = res.pts1[sl], res.pts2[sl]
        mine = gates(ofm.geometric_verification(p1, p2, res.F[p]), p1, p2)
        F_ref, _ = ofm.find_fundamental_cv2(p1, p2)
        theirs = gates(ofm.geometric_verification(p1, p2, F_ref), p1, p2) if F_ref is not None and np.asarray(F_ref).shape == (3, 3) else None
        if name in only_ours:      # the reference's run failed a gate: which one, and by how much?
            ok = theirs is None or failing_gate_is_marginal(theirs)
        else:                      # we fail a gate the reference's run passed
            ok = failing_gate_is_marginal(mine) or int(ref[name]["num_matches"]) != int(res.count[p])
        report.append((name, "ours only" if name in only_ours else "reference only", int(res.count[p]), mine, theirs, ok))
        if not ok:
            bad.append(name)
    with capsys.disabled():
        print(f"\npassing pairs: ours {len(ours)}, reference {len(ref)}, common {len(ours & set(ref))}")
        for r in report:
            print("  ", r)
    assert not bad, bad
    assert len(only_ours) + len(only_ref) <= 12


@pytest.mark.timeout(900)
def test_bench_36x8k_all_pairs_vs_cv2():
    """BASELINE configs[1]: the synthetic 36 x 8192 set of the bench (seed BASE_SEED + 1), all 630 pairs:
    kNN + ratio match lists (indices and DMatch.distance bits) equal to cv2's on every pair."""
    from sfm_b200 import lib as L, synth
    s = synth.make_image_set(36, 8192, seed=synth.BASE_SEED + 1, tie_frac=0.01)
    ctx = L.Context(0)
    ctx.set_images(s.desc.cuda(), s.kps.cuda(), s.offsets.numpy())
    pairs = synth.all_pairs(36).numpy()
    knn, ro = ctx.knn2(pairs)
    mq, mt, md2, cnt = ctx.filter_matches(pairs, ro, knn)
    torch.cuda.synchronize()
    mq, mt, md2, cnt = (t.cpu().numpy() for t in (mq, mt, md2, cnt))
    imgs = [s.image(i)[0].numpy() for i in range(36)]
    total = 0
    for p, (i, j) in enumerate(pairs):
        q, t, d = oknn.match_features_cv2(imgs[i], imgs[j])
        sl = slice(int(ro[p]), int(ro[p]) + int(cnt[p]))
        assert cnt[p] == len(q), (i, j, cnt[p], len(q))
        assert np.array_equal(mq[sl], q) and np.array_equal(mt[sl], t), (i, j)
        assert np.array_equal(np.sqrt(md2[sl].astype(np.float32)).view(np.uint32), d.view(np.uint32)), (i, j)
        total += len(q)
    assert total > 100000
    ctx.close()


@pytest.mark.timeout(600)
def test_ransac_envelope_synthetic_8k_vs_cv2():
    """SURVEY 8c criteria on 210 synthetic pairs of the bench's shape (median ~410 matches per pair, the
    chunk of the workload bunny does not reach), reference = cv2.findFundamentalMat + the reference's
    verification on the same correspondences.  Over the pairs the reference passes: (1) one-sided inlier
    bound on every pair, (2) Jaccard median >= 0.95, p10 >= 0.85, at most 2 pairs under 0.85, (4) no gate
    flip in either direction beyond 2 pairs; (3)/(5) mean error of our inliers <= 2 px, rank 2."""
    from sfm_b200 import lib as L, synth
    from sfm_b200.engine import PairEngine
    s = synth.make_image_set(21, 8192, seed=synth.BASE_SEED + 2, tie_frac=0.01)
    eng = PairEngine(0)
    eng.upload_images(s.desc, s.kps, s.offsets.numpy())
    pairs = synth.all_pairs(21).numpy()
    res = eng.match_pairs(pairs, only_ok=False)
    assert 300 <= np.median(res.count) <= 600
    jac, flips, low, n_big = [], 0, 0, 0
    for p in range(len(pairs)):
        sl = res.pair_slice(p)
        p1, p2 = res.pts1[sl], res.pts2[sl]
        if len(p1) < 8:
            continue
        F_ref, _ = ofm.find_fundamental_cv2(p1, p2)
        if F_ref is None or np.asarray(F_ref).shape != (3, 3):
            continue
        ref = ofm.geometric_verification(p1, p2, F_ref)
        ref_ok = bool(ofm.verify_match_quality(ref))
        ours_ok = res.status[p] == L.PAIR_OK
        flips += ref_ok != ours_ok
        if not ref_ok:
            continue
        n_big += len(p1) > 2048
        mine = res.inlier_mask[sl]
        v = ofm.geometric_verification(p1, p2, res.F[p])
        assert np.array_equal(v["inlier_mask"], mine)                       # K4 equals the oracle on OUR F
        I_new, I_ref, M = int(mine.sum()), int(ref["inlier_mask"].sum()), len(p1)
        assert I_new >= I_ref - max(3, 0.10 * M), (pairs[p], I_new, I_ref, M)
        j = (mine & ref["inlier_mask"]).sum() / max(1, (mine | ref["inlier_mask"]).sum())
        jac.append(j)
        low += M >= 60 and j < 0.85
        if ours_ok:
            assert v["metrics"]["reprojection_error"] <= 2.0
        sv = np.linalg.svd(res.F[p], compute_uv=False)
        assert sv[2] / sv[0] < 1e-6
    jac = np.array(jac)
    print(f"\nsynthetic 8k RANSAC parity: {len(jac)} reference-passing pairs, jaccard median {np.median(jac):.3f} "
          f"p10 {np.percentile(jac, 10):.3f} min {jac.min():.3f}, below 0.85: {low}, gate flips {flips}")
    assert len(jac) >= 120
    assert np.median(jac) >= 0.95 and np.percentile(jac, 10) >= 0.85
    assert low <= 2 and flips <= 2


def test_image_set_without_features():
    """Images without a single descriptor: the reference returns None per pair at :268; here every pair
    carries TOO_FEW_FEATURES and nothing raises (also through the bare kNN / filter entry points)."""
    from sfm_b200 import lib as L
    from sfm_b200.engine import PairEngine
    eng = PairEngine(0)
    eng.upload_images(torch.zeros((0, 128), dtype=torch.uint8), torch.zeros((0, 2), dtype=torch.float32),
                      np.zeros(4, np.int64))
    res = eng.match_pairs(np.array([(0, 1), (0, 2), (1, 2)], np.int32), only_ok=False)
    assert list(res.status) == [L.PAIR_TOO_FEW_FEATURES] * 3 and int(res.count.sum()) == 0
    knn, ro = eng.ctx.knn2([(0, 1)])
    assert knn.shape == (0, 4) and list(ro) == [0, 0]
    # one empty image among real ones: only its pairs are skipped
    rng = np.random.default_rng(3)
    d = torch.from_numpy(rng.integers(0, 200, (900, 128)).astype(np.uint8))
    k = torch.from_numpy((rng.random((900, 2)) * 700).astype(np.float32))
    eng.upload_images(d, k, np.array([0, 400, 400, 900], np.int64))
    res = eng.match_pairs(np.array([(0, 1), (1, 2), (0, 2)], np.int32), only_ok=False)
    assert res.status[0] == L.PAIR_TOO_FEW_FEATURES and res.status[1] == L.PAIR_TOO_FEW_FEATURES
    assert res.status[2] != L.PAIR_TOO_FEW_FEATURES
