"""Per-launch timing of the wide kNN kernel (B200 box; not a pytest module)."""
import json
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sfm_b200 import lib as L, synth  # noqa: E402

n_img = int(sys.argv[1]) if len(sys.argv) > 1 else 12
s = synth.make_image_set(n_img, 8192, seed=synth.BASE_SEED + 1, device="cuda", tie_frac=0.01)
ctx = L.Context(0)
ctx.set_images(s.desc, s.kps, s.offsets.numpy())
pairs = np.array([(i, j) for i in range(n_img) for j in range(i + 1, n_img)], np.int32)
ro = ctx.query_rows(pairs)
buf = ctx.empty((int(ro[-1]), 4), torch.int32)
ops = 256.0 * 8192 * 8192 * len(pairs)
out = {}
for mode, pf in ((5, 4), (5, 0), (6, 0), (5, 4)):
    ctx.set_knn_mode(mode)
    ctx.set_knn_probe_flags(pf)
    ts = []
    for i in range(4):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ctx.knn2(pairs, ro, buf)
        e1.record()
        torch.cuda.synchronize()
        ts.append(round(ops / (e0.elapsed_time(e1) * 1e-3) / 1e12, 1))
    out[f"mode{mode}_pf{pf}"] = ts
    out[f"brute_rows_after_mode{mode}_pf{pf}"] = ctx.brute_rows()
print(json.dumps(out))
