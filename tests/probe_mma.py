"""MMA-rate probes (B200 box; not a pytest module): tensor pipe alone at N = 64/128/256,
and the kNN kernel's TMA + MMA pipeline without an epilogue (mode 4) / fast path only (3)."""
import json
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sfm_b200 import lib as L, synth  # noqa: E402

ctx = L.Context(0)
out = {}
for n in (64, 128, 256):
    out[f"mma_only_n{n}_tops"] = round(ctx.probe_int8_rate(n, 4096)[0], 1)
for (k, fl, cons) in [(5, 3, 8), (5, 67, 8)]:
    out[f"hs_k{k}_f{fl}_c{cons}_tops"] = round(ctx.mma_handshake_probe(k, fl, cons), 1)
print(json.dumps(out))
n_img = 24
s = synth.make_image_set(n_img, 8192, seed=synth.BASE_SEED + 1, device="cuda", tie_frac=0.01)
ctx.set_images(s.desc, s.kps, s.offsets.numpy())
pairs = np.array([(i, j) for i in range(n_img) for j in range(i + 1, n_img)], np.int32)
ro = ctx.query_rows(pairs)
buf = ctx.empty((int(ro[-1]), 4), torch.int32)
ops = 256.0 * 8192 * 8192 * len(pairs)
for mode, pf in ((4, 0), (3, 0), (3, 2), (2, 0), (1, 0)):
    ctx.set_knn_mode(mode)
    ctx.set_knn_probe_flags(pf)
    for (w, e) in ([(0, 0), (2000, 200), (2000, 2000), (20000, 20000)] if mode == 2 else [(0, 0)]):
        ctx.set_knn_waits(w, e)
        ctx.knn2(pairs, ro, buf)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            ctx.knn2(pairs, ro, buf)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        out[f"mode{mode}_pf{pf}_w{w}_e{e}_tops"] = round(ops / (ms * 1e-3) / 1e12, 1)
print(json.dumps(out))
