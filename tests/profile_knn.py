"""Profiling driver (run under ncu on the GPU box): a few launches of the kNN
kernel on the 36 x 8192 workload, nothing else.  argv: n_images [knn_mode [probe_flags]]"""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sfm_b200 import lib as L, synth  # noqa: E402

n_img = int(sys.argv[1]) if len(sys.argv) > 1 else 36
mode = int(sys.argv[2]) if len(sys.argv) > 2 else 0
pf = int(sys.argv[3]) if len(sys.argv) > 3 else 0
s = synth.make_image_set(n_img, 8192, seed=synth.BASE_SEED + 1, device="cuda", tie_frac=0.01)
ctx = L.Context(0)
ctx.set_images(s.desc, s.kps, s.offsets.numpy())
ctx.set_knn_mode(mode)
ctx.set_knn_probe_flags(pf)
pairs = np.array([(i, j) for i in range(n_img) for j in range(i + 1, n_img)], np.int32)
ro = ctx.query_rows(pairs)
buf = ctx.empty((int(ro[-1]), 4), torch.int32)
for _ in range(3):
    ctx.knn2(pairs, ro, buf)
torch.cuda.synchronize()
print("ok")
