"""Same-box A/B of the kNN kernel variants (B200 box; not a pytest module):
mode 5 wide (2 x 256-column accumulators), 12 CTA pair (4 x 128), 15 wide + skip, 6/13 their max-tree-only
probes, 7/14 their no-epilogue probes.  Useful int8 TOP/s = 256 Nq Nt per pair / time."""
import json
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sfm_b200 import lib as L, synth  # noqa: E402

n_img = int(sys.argv[1]) if len(sys.argv) > 1 else 24
n_desc = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
modes = [int(m) for m in sys.argv[3].split(",")] if len(sys.argv) > 3 else [5, 12, 15, 6, 13, 7, 14, 5, 12]
waits = [tuple(int(x) for x in w.split(":")) for w in sys.argv[4].split(",")] if len(sys.argv) > 4 else [(0, 0)]
pflags = [int(x) for x in sys.argv[5].split(",")] if len(sys.argv) > 5 else [0]
s = synth.make_image_set(n_img, n_desc, seed=synth.BASE_SEED + 1, device="cuda", tie_frac=0.01)
ctx = L.Context(0)
ctx.set_images(s.desc, s.kps, s.offsets.numpy())
pairs = np.array([(i, j) for i in range(n_img) for j in range(i + 1, n_img)], np.int32)
ro = ctx.query_rows(pairs)
buf = ctx.empty((int(ro[-1]), 4), torch.int32)
ops = 256.0 * n_desc * n_desc * len(pairs)
out = {"pairs": len(pairs), "n_desc": n_desc}
ref = None
for mode, (w_ns, e_ns), pf in [(m, w, f) for m in modes for w in waits for f in pflags]:
    ctx.set_knn_mode(mode)
    ctx.set_knn_probe_flags(pf)
    ctx.set_knn_waits(w_ns, e_ns)
    ctx.knn2(pairs, ro, buf)
    torch.cuda.synchronize()
    ts = []
    for i in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ctx.knn2(pairs, ro, buf)
        e1.record()
        torch.cuda.synchronize()
        ts.append(round(ops / (e0.elapsed_time(e1) * 1e-3) / 1e12, 1))
    out.setdefault(f"mode{mode}_w{w_ns}_e{e_ns}_pf{pf}_tops", []).append(ts)
    if mode in (5, 12, 15, 16, 17) and pf == 0:
        got = buf.clone()
        if ref is None:
            ref = got
        out[f"mode{mode}_equals_first"] = bool(torch.equal(got, ref))
        out[f"brute_rows_mode{mode}"] = ctx.brute_rows()
    print(json.dumps(out), flush=True)
ctx.set_knn_mode(0)
ctx.set_knn_waits(0, 0)
ctx.set_knn_probe_flags(0)
Path("gpurun_out").mkdir(exist_ok=True)
Path(f"gpurun_out/r02_ab_knn_{n_img}x{n_desc}.json").write_text(json.dumps(out, indent=1))
