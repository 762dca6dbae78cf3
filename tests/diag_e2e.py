"""Where the end-to-end step spends its time (B200 box; not a pytest module)."""
import json
import sys
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sfm_b200 import synth  # noqa: E402
from sfm_b200.engine import PairEngine  # noqa: E402

n = 36
s = synth.make_image_set(n, 8192, seed=synth.BASE_SEED + 1, device="cuda", tie_frac=0.01)
off = s.offsets.numpy()
pairs = np.array([(i, j) for i in range(n) for j in range(i + 1, n)], np.int32)
eng = PairEngine(0)
h_desc = s.desc.cpu().pin_memory()
h_kps = s.kps.cpu().pin_memory()
dev = torch.device("cuda", 0)
out = {}


def t(tag, f, reps=5):
    f()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        r = f()
    torch.cuda.synchronize()
    out[tag] = round((time.perf_counter() - t0) / reps * 1e3, 3)
    return r


d = t("h2d", lambda: (h_desc.to(dev, non_blocking=True), h_kps.to(dev, non_blocking=True)))
t("set_images", lambda: eng.set_images(d[0], d[1], off))
t("match_device", lambda: eng.match_device(pairs))
t("match_pairs", lambda: eng.match_pairs(pairs))
ctx = eng.ctx
ro = ctx.query_rows(pairs)
knn = eng._buf("knn", (int(ro[-1]), 4), torch.int32)
t("query_rows", lambda: ctx.query_rows(pairs))
t("knn2", lambda: ctx.knn2(pairs, ro, knn))
print(json.dumps(out))
