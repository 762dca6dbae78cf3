"""TMEM read-rate probe + wait-flavour sweep of the kNN kernel (B200 box; not a pytest module)."""
import json
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sfm_b200 import lib as L, synth  # noqa: E402

ctx = L.Context(0)
out = {}
for warps in (4, 8, 16):
    for variant in (0, 1, 2, 3):
        ctx.tmem_ld_probe(warps, variant, 512)
        bpc, cyc = ctx.tmem_ld_probe(warps, variant, 4096)
        out[f"tmem_ld_w{warps}_v{variant}_B_per_cyc_per_SM"] = round(bpc, 1)
print(json.dumps(out))
n_img = 24
s = synth.make_image_set(n_img, 8192, seed=synth.BASE_SEED + 1, device="cuda", tie_frac=0.01)
ctx.set_images(s.desc, s.kps, s.offsets.numpy())
pairs = np.array([(i, j) for i in range(n_img) for j in range(i + 1, n_img)], np.int32)
ro = ctx.query_rows(pairs)
buf = ctx.empty((int(ro[-1]), 4), torch.int32)
ops = 256.0 * 8192 * 8192 * len(pairs)
res = {}
for mode in (3, 2):
    ctx.set_knn_mode(mode)
    for (w, e) in [(0, 0), (2000, 200), (2000, 0), (0, 200), (100, 50), (20000, 2000)]:
        ctx.set_knn_waits(w, e)
        ctx.knn2(pairs, ro, buf)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            ctx.knn2(pairs, ro, buf)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        res[f"mode{mode}_wait{w}_epi{e}_tops"] = round(ops / (ms * 1e-3) / 1e12, 1)
print(json.dumps(res))
