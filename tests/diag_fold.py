"""Diagnose the folded kNN path (B200 box; not a pytest module): ragged random sets vs the oracle,
and the brute-force row count on the 8k workload."""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sfm_b200 import lib as L, synth  # noqa: E402
from oracle import knn as oknn  # noqa: E402

ctx = L.Context(0)
rng = np.random.default_rng(5)
sizes = [1, 2, 7, 14, 15, 52, 127, 128, 129, 255, 256, 257, 300, 500, 513, 1000]
descs = [rng.integers(0, 256, (n, 128)).astype(np.uint8) for n in sizes]
off = np.zeros(len(sizes) + 1, np.int64)
off[1:] = np.cumsum(sizes)
ctx.set_images(torch.from_numpy(np.concatenate(descs)).cuda(), None, off)
print("flag", ctx.fold_flag())
pairs = [(i, j) for i in range(len(sizes)) for j in range(len(sizes)) if i != j and sizes[j] >= 2]
knn, ro = ctx.knn2(np.array(pairs, np.int32))
torch.cuda.synchronize()
got = knn.cpu().numpy()
print("exact-mode rows", ctx.brute_rows(), "of", len(got))
nbad = 0
for p, (i, j) in enumerate(pairs):
    idx, sqd = oknn.knn2_l2_numpy(descs[i], descs[j])
    g = got[ro[p]:ro[p + 1]]
    bad = np.flatnonzero((g[:, :2] != idx).any(1) | (g[:, 2:] != sqd).any(1))
    if len(bad):
        nbad += 1
        if nbad <= 12:
            r = bad[0]
            print("pair", (i, j), "sizes", sizes[i], sizes[j], "bad rows", len(bad), "first", r, "got", g[r], "want", idx[r], sqd[r])
print("bad pairs", nbad, "of", len(pairs))

n_img = 12
s = synth.make_image_set(n_img, 8192, seed=synth.BASE_SEED + 1, device="cuda", tie_frac=0.01)
ctx.set_images(s.desc, s.kps, s.offsets.numpy())
pairs = np.array([(i, j) for i in range(n_img) for j in range(i + 1, n_img)], np.int32)
knn, ro = ctx.knn2(pairs)
torch.cuda.synchronize()
print("8k: exact-mode rows", ctx.brute_rows(), "of", int(ro[-1]))
