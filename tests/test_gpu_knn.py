"""GPU parity of K1 (kNN k=2) and K2 (ratio / mutual filter) through the C ABI,
against the oracle and the golden vectors generated from cv2 (the reference's
own call, utils/find_matches.py:144-153 with NORM_L2)."""
import numpy as np
import pytest

from oracle import knn as oknn

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

INT32_MAX = 2**31 - 1


@pytest.fixture(scope="module")
def ctx():
    from sfm_b200 import lib as L
    c = L.Context(0)
    yield c
    c.close()


def register(ctx, descs, kps=None):
    counts = [len(d) for d in descs]
    off = np.zeros(len(descs) + 1, np.int64)
    off[1:] = np.cumsum(counts)
    cat = np.concatenate([np.asarray(d, np.uint8).reshape(-1, 128) for d in descs]) if off[-1] else np.zeros((0, 128), np.uint8)
    d = torch.from_numpy(cat).cuda()
    k = None
    if kps is not None:
        k = torch.from_numpy(np.concatenate(kps).astype(np.float32)).cuda()
    ctx.set_images(d, k, off)
    return off


def gpu_knn(ctx, pairs):
    knn, ro = ctx.knn2(np.asarray(pairs, np.int32))
    torch.cuda.synchronize()
    return knn.cpu().numpy(), ro


def check_pair(got, d1, d2):
    idx, sqd = oknn.knn2_l2_numpy(d1, d2)
    assert np.array_equal(got[:, :2], idx), "neighbour indices differ"
    assert np.array_equal(got[:, 2:], sqd), "squared distances differ"


@pytest.mark.parametrize("tag", ["rand", "ties", "dups", "tiny", "ragged"])
def test_knn_golden_cv2_vectors(ctx, golden_dir, tag):
    z = np.load(golden_dir / "knn_vectors.npz")
    d1, d2 = z[tag + "_d1"], z[tag + "_d2"]
    register(ctx, [d1, d2])
    got, ro = gpu_knn(ctx, [(0, 1)])
    assert np.array_equal(got[:, :2], z[tag + "_idx"])                       # bit-exact indices, cv2 tie rule
    dist = np.sqrt(got[:, 2:].astype(np.float32))
    assert np.array_equal(dist.view(np.uint32), z[tag + "_dist"].view(np.uint32))  # DMatch.distance bitwise
    mq, mt, md2, cnt = ctx.filter_matches([(0, 1)], ro, torch.from_numpy(got).cuda())
    n = int(cnt.cpu()[0])
    assert np.array_equal(mq.cpu().numpy()[:n], z[tag + "_good_q"])
    assert np.array_equal(mt.cpu().numpy()[:n], z[tag + "_good_t"])


def test_knn_ragged_sizes_vs_oracle(ctx):
    rng = np.random.default_rng(5)
    sizes = [1, 2, 7, 14, 15, 52, 127, 128, 129, 255, 256, 257, 300, 500, 513, 1000]
    descs = [rng.integers(0, 256, (n, 128)).astype(np.uint8) for n in sizes]
    off = register(ctx, descs)
    pairs = [(i, j) for i in range(len(sizes)) for j in range(len(sizes)) if i != j and sizes[j] >= 2]
    got, ro = gpu_knn(ctx, pairs)
    for p, (i, j) in enumerate(pairs):
        check_pair(got[ro[p]:ro[p + 1]], descs[i], descs[j])


def test_knn_tiny_train_images(ctx):
    rng = np.random.default_rng(6)
    descs = [rng.integers(0, 256, (n, 128)).astype(np.uint8) for n in (40, 1, 0, 3)]
    register(ctx, descs)
    got, ro = gpu_knn(ctx, [(0, 1), (0, 2), (3, 1), (1, 0), (2, 0)])
    a = got[ro[0]:ro[1]]                       # Nt = 1: best only
    assert np.all(a[:, 0] == 0) and np.all(a[:, 1] == -1) and np.all(a[:, 3] == INT32_MAX)
    want = ((descs[0].astype(np.int64) - descs[1][0].astype(np.int64)) ** 2).sum(1)
    assert np.array_equal(a[:, 2], want)
    b = got[ro[1]:ro[2]]                       # Nt = 0
    assert np.all(b[:, :2] == -1) and np.all(b[:, 2:] == INT32_MAX)
    assert ro[5] - ro[4] == 0                  # Nq = 0 -> no rows
    check_pair(got[ro[3]:ro[4]], descs[1], descs[0])


def test_knn_sift_like_ties_and_duplicates(ctx):
    from sfm_b200 import synth
    s = synth.make_image_set(4, [700, 900, 333, 1024], seed=77, tie_frac=0.05)
    descs = [s.image(i)[0].numpy() for i in range(4)]
    register(ctx, descs)
    pairs = [(i, j) for i in range(4) for j in range(4)]     # includes self pairs
    got, ro = gpu_knn(ctx, pairs)
    for p, (i, j) in enumerate(pairs):
        check_pair(got[ro[p]:ro[p + 1]], descs[i], descs[j])


def test_bunny_sift_all_pairs(ctx, golden_dir):
    """Real SIFT descriptors of the 36 bunny images (52..500 per image)."""
    z = np.load(golden_dir / "bunny_sift.npz")
    off = z["offsets"]
    descs = [z["desc"][off[i]:off[i + 1]] for i in range(len(off) - 1)]
    register(ctx, descs)
    n = len(descs)
    pairs = np.array([(i, j) for i in range(n) for j in range(i + 1, n)], np.int32)
    knn, ro = ctx.knn2(pairs)
    mq, mt, md2, cnt = ctx.filter_matches(pairs, ro, knn)
    torch.cuda.synchronize()
    got = knn.cpu().numpy()
    mq, mt, md2, cnt = (t.cpu().numpy() for t in (mq, mt, md2, cnt))
    for p, (i, j) in enumerate(pairs):
        check_pair(got[ro[p]:ro[p + 1]], descs[i], descs[j])
        q, t, d = oknn.match_features(descs[i], descs[j])
        s = slice(int(ro[p]), int(ro[p]) + int(cnt[p]))
        assert np.array_equal(mq[s], q) and np.array_equal(mt[s], t)
        assert np.array_equal(np.sqrt(md2[s].astype(np.float32)), d)


def test_mutual_nn_filter(ctx):
    rng = np.random.default_rng(11)
    d1 = rng.integers(0, 64, (300, 128)).astype(np.uint8)
    d2 = rng.integers(0, 64, (700, 128)).astype(np.uint8)
    d2[5] = d1[9]
    d2[600] = d1[9]          # duplicate train rows: lowest index must win both ways
    d1[17] = d1[9]
    register(ctx, [d1, d2])
    fwd, ro = ctx.knn2([(0, 1)])
    rev, ro_rev = ctx.knn2([(1, 0)])
    for ratio in (0.75, 2.0):            # 2.0 keeps every row: pure cross-check
        mq, mt, md2, cnt = ctx.filter_matches([(0, 1)], ro, fwd, ratio, ro_rev, rev)
        n = int(cnt.cpu()[0])
        q, t, d = oknn.match_features(d1, d2, ratio=ratio, mutual=True)
        assert np.array_equal(mq.cpu().numpy()[:n], q)
        assert np.array_equal(mt.cpu().numpy()[:n], t)
        assert np.array_equal(np.sqrt(md2.cpu().numpy()[:n].astype(np.float32)), d)


def test_knn_8k_config_sample_vs_oracle_and_properties(ctx):
    """BASELINE config 2 shape (8192-descriptor images): a sample of pairs against the
    oracle, then size-independent properties on all pairs of the small set."""
    from sfm_b200 import synth
    s = synth.make_image_set(6, 8192, seed=synth.BASE_SEED + 1, device="cuda", tie_frac=0.01)
    ctx.set_images(s.desc, s.kps, s.offsets.numpy())
    n = 6
    pairs = np.array([(i, j) for i in range(n) for j in range(n)], np.int32)
    knn, ro = ctx.knn2(pairs)
    knn2, _ = ctx.knn2(pairs)
    torch.cuda.synchronize()
    assert torch.equal(knn, knn2)                              # deterministic
    got = knn.cpu().numpy().reshape(n, n, 8192, 4)
    descs = [s.image(i)[0].cpu().numpy() for i in range(n)]
    for (i, j) in [(0, 1), (1, 0), (2, 5), (4, 4)]:
        check_pair(got[i, j], descs[i], descs[j])
    assert np.all(got[..., 2] <= got[..., 3])                  # ascending distances
    assert np.all(got[..., 0] != got[..., 1])
    for i in range(n):                                         # self match: distance 0
        assert np.all(got[i, i, :, 2] == 0)
        assert np.all(got[i, i, :, 0] <= np.arange(8192))      # lowest index among duplicates
    for i in range(n):                                         # min over rows == min over columns
        for j in range(i + 1, n):
            assert got[i, j, :, 2].min() == got[j, i, :, 2].min()


def test_knn_32k_config_sample_vs_oracle_and_properties(ctx):
    """BASELINE config 5 shape (32768-descriptor images, 128 train tiles per work item: 10-bit unit
    ids, truncated maxima in the keys): sampled query rows against the oracle, then
    size-independent properties over the full 32k x 32k pairs."""
    from sfm_b200 import synth
    n, nd = 3, 32768
    s = synth.make_image_set(n, nd, seed=synth.BASE_SEED + 4, device="cuda", tie_frac=0.01)
    ctx.set_images(s.desc, s.kps, s.offsets.numpy())
    pairs = np.array([(i, j) for i in range(n) for j in range(n)], np.int32)
    knn, ro = ctx.knn2(pairs)
    knn2, _ = ctx.knn2(pairs)
    torch.cuda.synchronize()
    assert torch.equal(knn, knn2)                              # deterministic
    got = knn.cpu().numpy().reshape(n, n, nd, 4)
    descs = [s.image(i)[0].cpu().numpy() for i in range(n)]
    rows = np.random.default_rng(5).choice(nd, 192, replace=False)
    rows.sort()
    for (i, j) in [(0, 1), (2, 0), (1, 1)]:
        idx, sqd = oknn.knn2_l2_numpy(descs[i][rows], descs[j])
        assert np.array_equal(got[i, j][rows, :2], idx), "neighbour indices differ"
        assert np.array_equal(got[i, j][rows, 2:], sqd), "squared distances differ"
    assert np.all(got[..., 2] <= got[..., 3])                  # ascending distances
    assert np.all(got[..., 0] != got[..., 1])
    assert np.all((got[..., :2] >= 0) & (got[..., :2] < nd))
    for i in range(n):                                         # self match: distance 0, lowest index
        assert np.all(got[i, i, :, 2] == 0)
        assert np.all(got[i, i, :, 0] <= np.arange(nd))
    for i in range(n):                                         # min over rows == min over columns
        for j in range(i + 1, n):
            assert got[i, j, :, 2].min() == got[j, i, :, 2].min()


# ---------------------------------------------------------------- kernel variants
FOLD_C = 255 * 31 * 255 + 254          # sfm_b200/csrc/knn_wide.cu kFoldC


def _sift_like(sizes, seed):
    from sfm_b200 import synth
    s = synth.make_image_set(len(sizes), list(sizes), seed=seed, tie_frac=0.03)
    return [s.image(i)[0].numpy() for i in range(len(sizes))]


def test_wide_fold_tile_accumulators(ctx):
    """Wide kernel path (N = 256 MMAs, fold operand as SWIZZLE_32B 32-byte rows): accumulators of a
    128 x 256 tile must equal a.b + C - (||b||^2 >> 1) exactly for every in-range row/column."""
    descs = _sift_like([300, 1500, 700], 31)
    off = register(ctx, descs)
    total = int(off[-1])
    allrows = np.concatenate(descs).astype(np.int64)
    for q_row, t_row in [(0, 300), (100, 300 + 512), (1800 + 131, 300 + 1024), (300 + 640, 1800), (7, 0),
                         (0, total - 200)]:
        got = ctx.debug_wide_tile(q_row, t_row).cpu().numpy()
        a = allrows[q_row:q_row + 128]
        b = allrows[t_row:t_row + 256]
        want = a @ b.T + FOLD_C - ((b * b).sum(1) >> 1)[None, :]
        assert np.array_equal(got[:len(a), :len(b)], want)


def test_pair_fold_tile_accumulators(ctx):
    """CTA-pair path (tcgen05.mma.cta_group::2, M = 256 over two CTAs, each staging its 128 query rows and
    64 of the 128 train rows): accumulators of a 256 x 128 tile must equal a.b + C - (||b||^2 >> 1) exactly."""
    descs = _sift_like([300, 1500, 700], 31)
    off = register(ctx, descs)
    total = int(off[-1])
    allrows = np.concatenate(descs).astype(np.int64)
    for q_row, t_row in [(0, 300), (100, 300 + 512), (1800 + 131, 300 + 1024), (300 + 640, 1800), (7, 0),
                         (total - 256, total - 100)]:
        got = ctx.debug_pair_tile(q_row, t_row).cpu().numpy()
        a = allrows[q_row:q_row + 256]
        b = allrows[t_row:t_row + 128]
        want = a @ b.T + FOLD_C - ((b * b).sum(1) >> 1)[None, :]
        assert np.array_equal(got[:len(a), :len(b)], want)


def test_ts_fold_tile_accumulators(ctx):
    """TS path (query rows written to tensor memory with tcgen05.st, tcgen05.mma with the A operand in
    TMEM, N = 128): accumulators of a 128 x 128 tile must equal a.b + C - (||b||^2 >> 1) exactly."""
    descs = _sift_like([300, 1500, 700], 31)
    off = register(ctx, descs)
    total = int(off[-1])
    allrows = np.concatenate(descs).astype(np.int64)
    for q_row, t_row in [(0, 300), (100, 300 + 512), (1800 + 131, 300 + 1024), (300 + 640, 1800), (7, 0),
                         (total - 128, total - 100)]:
        got = ctx.debug_ts_tile(q_row, t_row).cpu().numpy()
        a = allrows[q_row:q_row + 128]
        b = allrows[t_row:t_row + 128]
        want = a @ b.T + FOLD_C - ((b * b).sum(1) >> 1)[None, :]
        assert np.array_equal(got[:len(a), :len(b)], want)


@pytest.mark.parametrize("mode", [5, 12, 15, 16, 17])
def test_knn_fixup_paths(ctx, mode):
    """Data made of few distinct descriptors: unit maxima tie everywhere, so the folded kernels'
    tied-unit recompute (in the kernel or through the tie list, also with the list overflowing) and their
    brute-force fix-up pass (list driven, and the overflow scan with the list shrunk to 4 entries) produce
    the rows; all must equal the oracle incl. the tie rule."""
    rng = np.random.default_rng(77)
    base = _sift_like([6], 41)[0]
    descs = [base[rng.integers(0, 6, n)] for n in (300, 1100, 520)]
    register(ctx, descs)
    pairs = [(0, 1), (1, 2), (2, 0), (1, 0)]
    # 128: tied units recomputed in the kernel, 512: tie list of 2 entries (wide and ping-pong kernels only)
    for pf in ((0, 8) if mode == 12 else (0, 8, 128, 512, 512 | 8)):
        ctx.set_knn_mode(mode)
        ctx.set_knn_probe_flags(pf)
        try:
            before = ctx.brute_rows()
            got, ro = gpu_knn(ctx, pairs)
            assert ctx.brute_rows() > before          # the fix-up pass really ran
        finally:
            ctx.set_knn_mode(0)
            ctx.set_knn_probe_flags(0)
        for p, (i, j) in enumerate(pairs):
            check_pair(got[ro[p]:ro[p + 1]], descs[i], descs[j])


@pytest.mark.parametrize("mode", [5, 12, 15, 16, 17])
def test_knn_long_train_images(ctx, mode):
    """Folded kernels with long train images (wide: beyond 8192 rows, TS: beyond 16384): unit ids need
    more than 8 key bits, the unit maxima are truncated in the packed keys and ties on the truncated
    value are recomputed."""
    sizes = [700, 9000, 20000]
    descs = _sift_like(sizes, 35)
    register(ctx, descs)
    ctx.set_knn_mode(mode)
    try:
        pairs = [(0, 1), (0, 2), (1, 2)]
        got, ro = gpu_knn(ctx, pairs)
    finally:
        ctx.set_knn_mode(0)
    for p, (i, j) in enumerate(pairs):
        check_pair(got[ro[p]:ro[p + 1]], descs[i], descs[j])


@pytest.mark.parametrize("mode", [1, 5, 12, 15, 16, 17])
def test_knn_variants_match_oracle(ctx, mode):
    """Generic (1), N = 128 folded (2), wide folded (5) and TS (8) kernels, forced, on SIFT-like data
    with partial tiles and ties."""
    sizes = [1, 2, 5, 127, 128, 129, 255, 256, 257, 383, 512, 513, 640, 1025, 2000]
    descs = _sift_like(sizes, 32)
    register(ctx, descs)
    ctx.set_knn_mode(mode)
    try:
        pairs = [(i, j) for i in range(len(sizes)) for j in range(len(sizes)) if sizes[j] >= 2]
        got, ro = gpu_knn(ctx, pairs)
    finally:
        ctx.set_knn_mode(0)
    for p, (i, j) in enumerate(pairs):
        check_pair(got[ro[p]:ro[p + 1]], descs[i], descs[j])


def test_fold_range_limit_and_fallback(ctx):
    """||b||^2 = 2 C + 1 = 4,032,059 is the largest norm the fold digits encode; one more and
    sfm_set_images raises the flag so the generic kernel runs.  Results are exact either way."""
    rng = np.random.default_rng(33)
    top = np.zeros(128, np.uint8)
    top[:62] = 255
    top[62:66] = (22, 4, 3, 0)                                 # 62*255^2 + 484 + 16 + 9 = 4,032,059
    assert int((top.astype(np.int64) ** 2).sum()) == 2 * FOLD_C + 1
    d1 = rng.integers(0, 120, (200, 128)).astype(np.uint8)
    d1[:, 64:] = 0
    d2 = rng.integers(0, 140, (300, 128)).astype(np.uint8)
    d2[17] = top
    d2[211] = top                                              # duplicate of the extreme row: tie -> lowest index
    d1[5] = top
    d1[6] = 0                                                  # zero row: norm 0 -> digits of C itself
    d2[40] = 0
    register(ctx, [d1, d2])
    assert ctx.fold_flag() == 0
    got, ro = gpu_knn(ctx, [(0, 1), (1, 0)])
    check_pair(got[ro[0]:ro[1]], d1, d2)
    check_pair(got[ro[1]:ro[2]], d2, d1)
    over = top.copy()
    over[65] = 1                                               # 4,032,060 -> h = C + 1
    d2[100] = over
    register(ctx, [d1, d2])
    assert ctx.fold_flag() == 1
    got, ro = gpu_knn(ctx, [(0, 1), (1, 0)])
    check_pair(got[ro[0]:ro[1]], d1, d2)
    check_pair(got[ro[1]:ro[2]], d2, d1)


def test_knn_uniform_random_bytes_use_fold(ctx):
    """Uniform random bytes (||b||^2 ~ 2.8 M) are inside the fold range; forced fold == oracle."""
    rng = np.random.default_rng(34)
    descs = [rng.integers(0, 256, (n, 128)).astype(np.uint8) for n in (300, 700)]
    register(ctx, descs)
    assert ctx.fold_flag() == 0
    got, ro = gpu_knn(ctx, [(0, 1), (1, 0)])
    check_pair(got[ro[0]:ro[1]], descs[0], descs[1])
    check_pair(got[ro[1]:ro[2]], descs[1], descs[0])
