"""find_2d3d_matches: oracle vs the reference's own outputs (CPU), and the CUDA
radius matcher vs both (GPU).  Everything is bit-exact: indices, order, values, dtypes."""
import numpy as np
import pytest

from oracle import tracks as otr


def load_cases(golden_dir):
    z = np.load(golden_dir / "find_2d3d_vectors.npz")
    cases = []
    for c in range(int(z["n_cases"])):
        pre = f"c{c}_"
        n_tracks = int(z[pre + "n_tracks"])
        tracks = [dict() for _ in range(n_tracks)]
        for k, i, x, y in z[pre + "tracks"]:
            tracks[int(k)][int(i)] = [float(x), float(y)]
        pairs = {}
        for key in z.files:
            if key.startswith(pre + "pair_") and key.endswith("_pts1"):
                name = key[len(pre):-5]
                pairs[name] = (z[key], z[pre + name + "_pts2"])
        dropped = str(z[pre + "dropped"])

        def load_pair(name, pairs=pairs, dropped=dropped):
            if name == dropped or name not in pairs:
                raise FileNotFoundError(name)
            return pairs[name]
        cases.append(dict(new_id=int(z[pre + "new_id"]), tracks=tracks, points3D=list(z[pre + "points3D"]),
                          order=[str(s) for s in z[pre + "order"]], load_pair=load_pair,
                          out3=z[pre + "out3"], out2=z[pre + "out2"]))
    return cases


def same(a, b):
    return a.shape == b.shape and a.dtype == b.dtype and np.array_equal(a, b)


# ----------------------------------------------------------------------------- CPU
def test_oracle_reproduces_reference_outputs(golden_dir):
    cases = load_cases(golden_dir)
    assert len(cases) >= 5 and sum(len(c["out3"]) for c in cases) > 500
    for c in cases:
        p3, p2 = otr.find_2d3d_matches(c["tracks"], c["points3D"], c["new_id"], c["order"], c["load_pair"])
        assert same(p3, c["out3"]) and same(p2, c["out2"])


def test_oracle_loops_equal_numpy():
    rng = np.random.default_rng(5)
    t = rng.uniform(0, 50, (40, 2)).astype(np.float32).astype(np.float64)
    p = np.concatenate([(t[:20] + [2.0, 0.0]).astype(np.float32), (t[20:] + [1.2, 1.6]).astype(np.float32),
                        rng.uniform(0, 50, (60, 2)).astype(np.float32)])
    a = otr.radius_match_numpy(t, p)
    b = otr.radius_match_loops(t, p)
    assert len(a[0]) > 10 and np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


def test_segments_skip_what_the_reference_skips(golden_dir):
    from sfm_b200 import tracks as T
    c = load_cases(golden_dir)[1]                      # has a missing correspondence file
    segs = T.build_segments(c["tracks"], c["new_id"], c["order"] + ["pair_x_y", "notapair"], c["load_pair"])
    assert 0 < len(segs) < len(c["order"]) + 2
    for idx, tp, other, new in segs:
        assert tp.dtype == np.float64 and tp.shape == (len(idx), 2) and len(other) == len(new)


# ----------------------------------------------------------------------------- GPU
@pytest.fixture(scope="module")
def ctx():
    from sfm_b200 import lib as L
    c = L.Context(0)
    yield c
    c.close()


def gpu_radius(ctx, segs, thr):
    import torch
    toff = np.cumsum([0] + [len(t) for t, _ in segs])
    poff = np.cumsum([0] + [len(p) for _, p in segs])
    tr = np.concatenate([np.asarray(t, np.float64).reshape(-1, 2) for t, _ in segs])
    pt = np.concatenate([np.asarray(p).astype(np.float64).reshape(-1, 2) for _, p in segs])
    so, ti, pi = ctx.radius_match(toff, torch.from_numpy(tr).cuda(), poff, torch.from_numpy(pt).cuda(), thr)
    return so, ti.cpu().numpy(), pi.cpu().numpy()


@pytest.mark.gpu
def test_gpu_find_2d3d_matches_equals_reference(ctx, golden_dir):
    from sfm_b200 import tracks as T
    for c in load_cases(golden_dir):
        p3, p2 = T.find_2d3d_matches(ctx, c["tracks"], c["points3D"], c["new_id"], c["order"], c["load_pair"])
        assert same(p3, c["out3"]) and same(p2, c["out2"])


@pytest.mark.gpu
@pytest.mark.parametrize("thr", [2.0, 0.5, 3.7, 1e-3, 0.0, -1.0, float("inf"), 1.9999999999999998, 2.0000000000000004])
def test_gpu_radius_match_threshold_edges(ctx, thr):
    """Points planted at f64 distances of exactly 2, one ulp around it, 0, and random ones; ragged
    segments incl. empty ones; every threshold decision must equal numpy's sqrt(dx^2 + dy^2) < thr."""
    rng = np.random.default_rng(11)
    segs = []
    for nt, npnt in [(300, 700), (0, 10), (5, 0), (1, 1), (257, 1025), (1000, 33)]:
        t = rng.uniform(0, 64, (nt, 2)).astype(np.float32).astype(np.float64)
        p = rng.uniform(0, 64, (npnt, 2)).astype(np.float32)
        if nt and npnt:
            k = rng.integers(0, nt, npnt // 2)
            offs = np.array([(2.0, 0.0), (0.0, 2.0), (1.2, 1.6), (0.0, 0.0), (np.nextafter(2.0, 3), 0.0),
                             (np.nextafter(2.0, 1), 0.0), (1.4142135623730951, 1.4142135623730951)])
            p = p.astype(np.float64)
            p[:len(k)] = t[k] + offs[rng.integers(0, len(offs), len(k))]     # f64 points: exact edges
        segs.append((t, p))
    so, ti, pi = gpu_radius(ctx, segs, thr)
    n_hits = 0
    for s, (t, p) in enumerate(segs):
        wt, wp = otr.radius_match_numpy(t, p, thr)
        a, b = so[s], so[s + 1]
        assert b - a == len(wt)
        assert np.array_equal(ti[a:b], wt) and np.array_equal(pi[a:b], wp)
        n_hits += len(wt)
    if thr == 2.0:
        assert n_hits > 300
    if thr <= 0:
        assert n_hits == 0
    if thr == float("inf"):
        assert n_hits == sum(len(t) * len(p) for t, p in segs)


@pytest.mark.gpu
def test_gpu_radius_match_large_vs_oracle_blocks(ctx):
    """12,000 tracks x 30,000 points in 3 segments (3.6e8 distance tests); the oracle runs in blocks of
    tracks.  Also the symmetry property: swapping the roles transposes the hit list."""
    rng = np.random.default_rng(12)
    segs = []
    for nt, npnt in [(6000, 30000), (4000, 20000), (2000, 5000)]:
        t = rng.uniform(0, [1024, 768], (nt, 2)).astype(np.float32).astype(np.float64)
        p = rng.uniform(0, [1024, 768], (npnt, 2)).astype(np.float32)
        segs.append((t, p))
    so, ti, pi = gpu_radius(ctx, segs, 2.0)
    for s, (t, p) in enumerate(segs):
        wt, wp = [], []
        for r0 in range(0, len(t), 500):
            a, b = otr.radius_match_numpy(t[r0:r0 + 500], p, 2.0)
            wt.append(a + r0)
            wp.append(b)
        wt, wp = np.concatenate(wt), np.concatenate(wp)
        a, b = so[s], so[s + 1]
        assert len(wt) > 100 and b - a == len(wt)
        assert np.array_equal(ti[a:b], wt) and np.array_equal(pi[a:b], wp)
    t, p = segs[2]
    so2, ti2, pi2 = gpu_radius(ctx, [(p.astype(np.float64), t)], 2.0)
    fwd = set(zip(ti[so[2]:so[3]].tolist(), pi[so[2]:so[3]].tolist()))
    assert fwd == set(zip(pi2.tolist(), ti2.tolist()))
