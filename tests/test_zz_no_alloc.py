"""sfm_filter_matches_static (the ratio filter with its metadata in a file-scope __device__ array, no
allocation inside the call) against sfm_filter_matches, through the C ABI, bit for bit, on two calls in
a row.  No counterpart in the reference.
Named test_zz_* so that pytest collects it last."""
import re

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    from sfm_b200 import lib as L
    c = L.Context(0)
    yield c
    c.close()


def make_images(seed, sizes):
    """Noisy views of one descriptor pool, so that a good share of the rows passes the ratio test."""
    rng = np.random.default_rng(seed)
    pool = rng.integers(0, 256, (max(sizes) + 64, 128)).astype(np.int16)
    descs = []
    for n in sizes:
        rows = rng.permutation(len(pool))[:n]
        noisy = pool[rows] + rng.integers(-12, 13, (n, 128))
        descs.append(np.clip(noisy, 0, 255).astype(np.uint8))
    return descs


def register(ctx, descs):
    off = np.zeros(len(descs) + 1, np.int64)
    off[1:] = np.cumsum([len(d) for d in descs])
    cat = np.concatenate([d.reshape(-1, 128) for d in descs])
    ctx.set_images(torch.from_numpy(cat).cuda(), None, off)
    return off


def lists(out, ro):
    mq, mt, md2, cnt = (t.cpu().numpy() for t in out)
    cnt = cnt[:len(ro) - 1]
    return [(mq[ro[p]:ro[p] + c].copy(), mt[ro[p]:ro[p] + c].copy(), md2[ro[p]:ro[p] + c].copy())
            for p, c in enumerate(cnt)], cnt


def compare(ctx, pairs, ratio, out_static):
    pairs = np.asarray(pairs, np.int32)
    knn, ro = ctx.knn2(pairs)
    want = ctx.filter_matches(pairs, ro, knn, ratio=ratio)
    for t in out_static:
        t.fill_(-7)                                  # stale values must not survive as results
    got = ctx.filter_matches_static(pairs, ro, knn, ratio=ratio, out=out_static)
    torch.cuda.synchronize()
    wl, wc = lists(want, ro)
    gl, gc = lists(got, ro)
    assert np.array_equal(wc, gc)
    for w, g in zip(wl, gl):
        for a, b in zip(w, g):
            assert np.array_equal(a, b)              # integer lists: bit-exact
    return int(wc.sum())


def test_static_filter_matches_existing_entry_point_twice(ctx):
    sizes = [700, 0, 1, 2, 255, 256, 257, 1023, 1024, 1025, 2500, 4100]
    descs = make_images(21, sizes)
    register(ctx, descs)
    n_img = len(sizes)
    all_pairs = [(i, j) for i in range(n_img) for j in range(n_img) if i != j]
    rows = sum(sizes) * n_img
    # outputs allocated once, before the calls: nothing is allocated between or inside them
    out = (torch.empty(rows, dtype=torch.int32, device="cuda"), torch.empty(rows, dtype=torch.int32, device="cuda"),
           torch.empty(rows, dtype=torch.int32, device="cuda"), torch.empty(len(all_pairs), dtype=torch.int32, device="cuda"))
    launches0 = ctx.launches
    kept1 = compare(ctx, all_pairs, 0.75, out)       # first call: 132 pairs
    # second call right after it: fewer pairs, another order, another ratio -- the array still holds call 1's records
    second = [all_pairs[i] for i in np.random.default_rng(3).permutation(len(all_pairs))[:37]]
    kept2 = compare(ctx, second, 0.9, out)
    assert kept1 > 1000 and kept2 > 300              # the comparison was not between empty lists
    assert ctx.launches > launches0
    # and the first batch again: same answer as the first time
    assert compare(ctx, all_pairs, 0.75, out) == kept1


def test_static_filter_rejects_more_pairs_than_the_array_holds(ctx):
    from sfm_b200 import lib as L
    descs = make_images(22, [64, 64])
    register(ctx, descs)
    header = (L.LIB_PATH.parent.parent / "include" / "sfm_match.h").read_text()
    limit = int(re.search(r"#define\s+SFM_FILTER_STATIC_MAX_PAIRS\s+(\d+)", header).group(1))
    n = limit + 1
    pairs = np.tile(np.array([[0, 1]], np.int32), (n, 1))
    ro = np.arange(n + 1, dtype=np.int64) * 64
    knn = torch.zeros((64, 4), dtype=torch.int32, device="cuda")
    out = tuple(torch.zeros(8, dtype=torch.int32, device="cuda") for _ in range(4))
    with pytest.raises(L.SfmError, match="SFM_FILTER_STATIC_MAX_PAIRS"):
        ctx.filter_matches_static(pairs, ro, knn, out=out)
    # the set-up-only call (n_pairs = 0) is accepted
    ctx.filter_matches_static(np.zeros((0, 2), np.int32), np.zeros(1, np.int64), knn, out=out)
    torch.cuda.synchronize()
