"""Small inputs through every hand-written kernel, for compute-sanitizer (B200 box):
smoke() (kNN wide + fix-up, filter, RANSAC + verification, pack, radius matcher), then the kNN kernels on a
ragged set with partial train tiles and two-tile images in the wide (5) and CTA-pair (12) variants, the
mutual filter and an all-ties set that drives the fix-up pass."""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import __graft_entry__ as g  # noqa: E402
from oracle import knn as oknn  # noqa: E402
from sfm_b200 import lib as L, synth  # noqa: E402

g.smoke()
s = synth.make_image_set(5, [300, 513, 40, 700, 129], seed=3, tie_frac=0.05)
ctx = L.Context(0)
ctx.set_images(s.desc.cuda(), s.kps.cuda(), s.offsets.numpy())
pairs = np.array([(0, 1), (1, 3), (3, 0), (2, 4), (4, 1), (1, 0)], np.int32)
for mode in (5, 12, 16, 17):
    ctx.set_knn_mode(mode)
    knn, ro = ctx.knn2(pairs)
    torch.cuda.synchronize()
    got = knn.cpu().numpy()
    for p, (i, j) in enumerate(pairs):
        idx, sqd = oknn.knn2_l2_numpy(s.image(int(i))[0].numpy(), s.image(int(j))[0].numpy())
        assert np.array_equal(got[ro[p]:ro[p + 1], :2], idx) and np.array_equal(got[ro[p]:ro[p + 1], 2:], sqd), (mode, i, j)
ctx.set_knn_mode(0)
base = s.image(0)[0][:5]
ties = torch.cat([base[torch.randint(0, 5, (n,))] for n in (260, 300)])
ctx.set_images(ties.cuda(), None, np.array([0, 260, 560], np.int64))
for mode in (5, 12, 16, 17):
    ctx.set_knn_mode(mode)
    knn, ro = ctx.knn2([(0, 1), (1, 0)])
    rev, ro_rev = ctx.knn2([(1, 0), (0, 1)])
    ctx.filter_matches([(0, 1), (1, 0)], ro, knn, 2.0, ro_rev, rev)
    torch.cuda.synchronize()
ctx.set_knn_mode(0)
print("sanitize_small ok, launches", ctx.launches)
