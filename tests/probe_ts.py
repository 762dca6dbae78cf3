"""TS-mode (A operand in TMEM) probes (B200 box; not a pytest module): the folded accumulators of
a 128 x 128 tile through tcgen05.st + TS-form MMAs against numpy, then issue rates of 128 x N x 32
u8 MMAs with A in TMEM vs in shared memory."""
import json
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sfm_b200 import lib as L, synth  # noqa: E402

ctx = L.Context(0)
s = synth.make_image_set(2, [640, 700], seed=5, device="cuda")
ctx.set_images(s.desc, s.kps, s.offsets.numpy())
d = s.desc.cpu().numpy().astype(np.int64)
out = {}
C = 255 * 31 * 255 + 254
for (q, t) in ((0, 640), (128, 640 + 256), (512, 0), (37, 640 + 131)):
    got = ctx.debug_ts_tile(q, t).cpu().numpy()
    a, b = d[q:q + 128], d[t:t + 128]
    want = a @ b.T + (C - (b * b).sum(1) // 2)[None, :]
    n = min(128, len(d) - t)
    out[f"tile_q{q}_t{t}_exact"] = bool(np.array_equal(got[:, :n], want[:, :n]))
    if not out[f"tile_q{q}_t{t}_exact"]:
        out[f"tile_q{q}_t{t}_sample"] = [got[:2, :4].tolist(), want[:2, :4].tolist()]
print(json.dumps(out), flush=True)
for n in (64, 96, 128, 192, 256):
    for ts in (1, 0):
        for commit in (0, 1):
            out[f"rate_n{n}_{'ts' if ts else 'ss'}_c{commit}"] = round(ctx.ts_rate_probe(n, ts, commit, 8192), 1)
print(json.dumps(out), flush=True)
Path("gpurun_out").mkdir(exist_ok=True)
Path("gpurun_out/r02_probe_ts.json").write_text(json.dumps(out, indent=1))
