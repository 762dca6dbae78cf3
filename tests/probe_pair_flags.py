"""Pair kernel, no-epilogue mode (14) under probe flags: which part of the pipeline limits the MMA rate."""
import json, sys
from pathlib import Path
import numpy as np, torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sfm_b200 import lib as L, synth  # noqa: E402
n_img = int(sys.argv[2]) if len(sys.argv) > 2 else 12
s = synth.make_image_set(n_img, 8192, seed=synth.BASE_SEED + 1, device="cuda", tie_frac=0.01)
ctx = L.Context(0)
ctx.set_images(s.desc, s.kps, s.offsets.numpy())
pairs = np.array([(i, j) for i in range(n_img) for j in range(i + 1, n_img)], np.int32)
ro = ctx.query_rows(pairs)
buf = ctx.empty((int(ro[-1]), 4), torch.int32)
ops = 256.0 * 8192 * 8192 * len(pairs)
out = {}
cfgs = [tuple(int(x) for x in c.split(":")) for c in sys.argv[1].split(",")]
for mode, pf in cfgs:
    ctx.set_knn_mode(mode)
    ctx.set_knn_probe_flags(pf)
    ctx.knn2(pairs, ro, buf)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    ctx.knn2(pairs, ro, buf)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    out[f"mode{mode}_pf{pf}"] = {"tops": round(ops / (ms * 1e-3) / 1e12, 1), "ms": round(ms, 3)}
    if pf & 512:
        c = ctx.pair_counters()
        n = max(c[5], 1)
        out[f"mode{mode}_pf{pf}"]["epi_per_item"] = {"main_loop": c[0] // n, "resolve": c[1] // n, "recompute": c[2] // n,
                                                    "merge_write": c[3] // n, "recomputed_rows_per_item_warp": round(c[4] / n, 2), "items": n}
    if pf & 256:
        c = buf.view(-1)[:16].cpu().numpy().view(np.int64)
        g = max(int(c[6]), 1)
        out[f"mode{mode}_pf{pf}"]["per_group"] = {k: round(float(c[i]) / g, 1) for i, k in
            enumerate(["total", "wait_b_full", "wait_t_empty", "mma_issue", "commit_t_full", "commit_b_empty"])}
        out[f"mode{mode}_pf{pf}"]["groups"] = g
    print(json.dumps(out), flush=True)
ctx.set_knn_mode(0); ctx.set_knn_probe_flags(0)
Path("gpurun_out/r02_probe_pair_flags.json").write_text(json.dumps(out, indent=1))
