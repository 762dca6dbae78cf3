"""RANSAC + verification kernel timing on the 36 x 8192 workload under different parameters
(B200 box; not a pytest module): shows what the hypothesis rounds / refits / verification cost."""
import json
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sfm_b200 import lib as L, synth  # noqa: E402

n_img = int(sys.argv[1]) if len(sys.argv) > 1 else 36
s = synth.make_image_set(n_img, 8192, seed=synth.BASE_SEED + 1, device="cuda", tie_frac=0.01)
ctx = L.Context(0)
ctx.set_images(s.desc, s.kps, s.offsets.numpy())
pairs = np.array([(i, j) for i in range(n_img) for j in range(i + 1, n_img)], np.int32)
ro = ctx.query_rows(pairs)
knn = ctx.empty((int(ro[-1]), 4), torch.int32)
ctx.knn2(pairs, ro, knn)
mq, mt, md2, cnt = ctx.filter_matches(pairs, ro, knn, 0.75, None, None)
c = cnt.cpu().numpy()
out = {"matches_per_pair": {"min": int(c.min()), "median": float(np.median(c)), "mean": float(c.mean()),
                            "max": int(c.max()), "lt8": int((c < 8).sum())}}


def run(tag, **kw):
    prm = L.default_params(**kw)
    for _ in range(3):
        v = ctx.verify_pairs(pairs, ro, mq, mt, cnt, prm)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        v = ctx.verify_pairs(pairs, ro, mq, mt, cnt, prm)
    e1.record()
    torch.cuda.synchronize()
    st = v["status"].cpu().numpy()
    out.setdefault(tag, {"ms": [], "ok": int((st == L.PAIR_OK).sum())})["ms"].append(round(e0.elapsed_time(e1) / 20, 3))


VARIANTS = [("default", {}), ("lo0", dict(lo_iters=0)), ("lo1", dict(lo_iters=1)),
            ("iters128", dict(max_iters=128, min_work=0)), ("iters128_lo0", dict(max_iters=128, min_work=0, lo_iters=0)),
            ("iters32", dict(max_iters=32, min_work=0)), ("iters32_lo0", dict(max_iters=32, min_work=0, lo_iters=0)),
            ("minwork0", dict(min_work=0))]
for rep in range(3):                       # interleaved: clock drift shows up as spread within a tag, not between tags
    for tag, kw in (VARIANTS if rep % 2 == 0 else VARIANTS[::-1]):
        run(tag, **kw)
print(json.dumps(out))
