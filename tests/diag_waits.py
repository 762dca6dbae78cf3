"""Wide kNN kernel under different mbarrier suspend hints (B200 box; not a pytest module)."""
import json
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sfm_b200 import lib as L, synth  # noqa: E402

n_img = 36
s = synth.make_image_set(n_img, 8192, seed=synth.BASE_SEED + 1, device="cuda", tie_frac=0.01)
ctx = L.Context(0)
ctx.set_images(s.desc, s.kps, s.offsets.numpy())
pairs = np.array([(i, j) for i in range(n_img) for j in range(i + 1, n_img)], np.int32)
ro = ctx.query_rows(pairs)
buf = ctx.empty((int(ro[-1]), 4), torch.int32)
ops = 256.0 * 8192 * 8192 * len(pairs)
out = {}
for w, e in ((0, 0), (0, 100), (0, 300), (0, 1000), (200, 0), (1000, 0), (1000, 300), (5000, 1000)):
    ctx.set_knn_waits(w, e)
    ctx.knn2(pairs, ro, buf)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(4):
        ctx.knn2(pairs, ro, buf)
    e1.record()
    torch.cuda.synchronize()
    out[f"w{w}_e{e}"] = round(ops / (e0.elapsed_time(e1) / 4 * 1e-3) / 1e12, 1)
print(json.dumps(out))
