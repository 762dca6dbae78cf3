"""Diagnose the folded kNN path on the golden 'ties' vectors (B200 box; not a pytest module)."""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sfm_b200 import lib as L  # noqa: E402
from oracle import knn as oknn  # noqa: E402

C = 255 * 31 * 255 + 254
z = np.load(Path(__file__).parent / "golden" / "knn_vectors.npz")
tag = sys.argv[1] if len(sys.argv) > 1 else "ties"
d1, d2 = z[tag + "_d1"].astype(np.uint8), z[tag + "_d2"].astype(np.uint8)
ctx = L.Context(0)
off = np.array([0, len(d1), len(d1) + len(d2)], np.int64)
ctx.set_images(torch.from_numpy(np.concatenate([d1, d2])).cuda(), None, off)
print("flag", ctx.fold_flag(), d1.shape, d2.shape)
b64 = d2.astype(np.int64)
nb = (b64 * b64).sum(1)
for q0 in (0, 128, 256):
    for tile in range((len(d2) + 127) // 128):
        got = ctx.debug_fold_tile(q0, 1, tile).cpu().numpy()
        a = d1[q0:q0 + 128].astype(np.int64)
        b = b64[tile * 128:(tile + 1) * 128]
        want = a @ b.T + C - (nb[tile * 128:(tile + 1) * 128] >> 1)[None, :]
        g = got[:len(a), :len(b)]
        bad = g != want
        print("q0", q0, "tile", tile, "mismatch", int(bad.sum()), "of", bad.size)
        if bad.any():
            r, c = np.argwhere(bad)[0]
            print("  first", r, c, "got", g[r, c], "want", want[r, c], "diff", g[r, c] - want[r, c],
                  "cols bad", np.flatnonzero(bad.any(0))[:10], "rows bad", np.flatnonzero(bad.any(1))[:10])
for mode in (1, 2, 0):
    ctx.set_knn_mode(mode)
    knn, ro = ctx.knn2(np.array([(0, 1)], np.int32))
    torch.cuda.synchronize()
    got = knn.cpu().numpy()
    idx, sqd = oknn.knn2_l2_numpy(d1, d2)
    bi = np.flatnonzero((got[:, :2] != idx).any(1))
    bd = np.flatnonzero((got[:, 2:] != sqd).any(1))
    print("mode", mode, "idx-bad rows", len(bi), "dist-bad rows", len(bd))
    for r in bi[:5]:
        print("  row", r, "got", got[r], "want", idx[r], sqd[r])
