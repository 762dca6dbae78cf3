"""Kernel-only timing of the kNN variants on the 36 x 8192 workload (run on the B200 box;
not a pytest module).  mode 1 generic, 2 folded, 3 folded fast path only (ceiling of the
epilogue design: TMEM reads + max tree, no exact path)."""
import json
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sfm_b200 import lib as L, synth  # noqa: E402

n_img = int(sys.argv[1]) if len(sys.argv) > 1 else 36
modes = [int(m) for m in sys.argv[2].split(",")] if len(sys.argv) > 2 else [2, 5, 6, 7]
s = synth.make_image_set(n_img, 8192, seed=synth.BASE_SEED + 1, device="cuda", tie_frac=0.01)
ctx = L.Context(0)
ctx.set_images(s.desc, s.kps, s.offsets.numpy())
pairs = np.array([(i, j) for i in range(n_img) for j in range(i + 1, n_img)], np.int32)
ro = ctx.query_rows(pairs)
buf = ctx.empty((int(ro[-1]), 4), torch.int32)
ops = 256.0 * 8192 * 8192 * len(pairs)
out = {"fold_flag": ctx.fold_flag()}
ref = None
for mode in modes:
    ctx.set_knn_mode(mode)
    for _ in range(2):
        ctx.knn2(pairs, ro, buf)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        ctx.knn2(pairs, ro, buf)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    out[f"mode{mode}_ms"] = ms
    out[f"mode{mode}_tops"] = ops / (ms * 1e-3) / 1e12
    if mode in (1, 2, 5):
        if ref is None:
            ref = buf.clone()
        else:
            out[f"mode{mode}_equals_first"] = bool(torch.equal(ref, buf))
ctx.set_knn_mode(5)
ctx.set_knn_probe_flags(4)
for _ in range(2):
    ctx.knn2(pairs, ro, buf)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    ctx.knn2(pairs, ro, buf)
e1.record()
torch.cuda.synchronize()
out["mode5_nofix_tops"] = ops / (e0.elapsed_time(e1) / 5 * 1e-3) / 1e12
out["flagged_rows_frac"] = float((buf[:, 0] == -2).float().mean().item())
ctx.set_knn_probe_flags(0)
print(json.dumps(out))
