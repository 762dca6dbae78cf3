"""Cycle breakdown of the TS kNN kernel (B200 box; not a pytest module)."""
import json
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sfm_b200 import lib as L, synth  # noqa: E402

n_img = int(sys.argv[1]) if len(sys.argv) > 1 else 16
s = synth.make_image_set(n_img, 8192, seed=synth.BASE_SEED + 1, device="cuda", tie_frac=0.01)
ctx = L.Context(0)
ctx.set_images(s.desc, s.kps, s.offsets.numpy())
pairs = np.array([(i, j) for i in range(n_img) for j in range(i + 1, n_img)], np.int32)
ro = ctx.query_rows(pairs)
buf = ctx.empty((int(ro[-1]), 4), torch.int32)
names = {6: "mma_commit_to_self_sees_full", 0: "mma_total", 1: "mma_wait_b_full", 2: "mma_wait_t_empty", 3: "mma_wait_a_full", 4: "prod_wait_b_empty",
         5: "groups", 8: "epi0_total", 9: "epi0_wait_t_full", 10: "epi0_process_group", 11: "epi0_item_end",
         12: "epi0_a_refill"}
out = {}
cfgs = [tuple(int(x) for x in c.split(":")) for c in sys.argv[2].split(",")] if len(sys.argv) > 2 else [(8, 0), (8, 16), (10, 16), (11, 16)]
for mode, pf in cfgs:
    ctx.set_knn_mode(mode)
    ctx.set_knn_probe_flags(pf)
    ctx.knn_profile(True)
    ctx.knn2(pairs, ro, buf)
    torch.cuda.synchronize()
    c = ctx.knn_profile(False)
    ev = ctx.knn_profile_events()
    g = np.arange(200, 3000)             # steady state, CTA 0
    even = g[g % 2 == 0]
    lat = {
        "commit_to_full_seen": float(np.mean(ev[2][g] - ev[1][g])),
        "full_seen_to_release": float(np.mean(ev[3][g] - ev[2][g])),
        "release_to_mma_passes_t_empty_of_g_plus_3": float(np.mean(ev[0][g + 3] - ev[3][g])),
        "mma_passes_t_empty_to_commit": float(np.mean(ev[1][g] - ev[0][g])),
        "group_period": float(np.mean(ev[1][g + 1] - ev[1][g])),
        "full_seen_to_group_done": float(np.mean(ev[4][g] - ev[2][g])),
        "commit_g_to_commit_g_plus_3": float(np.mean(ev[1][g + 3] - ev[1][g])),
        "blk0_group_done_to_next_full_seen": float(np.mean(ev[2][even + 2] - ev[4][even])),
    }
    groups = max(c[5], 1)
    out[f"mode{mode}_pf{pf}"] = {names[i]: round(c[i] / groups, 1) for i in names if i != 5}
    out[f"mode{mode}_pf{pf}"]["groups_per_cta"] = groups / 148
    out[f"mode{mode}_pf{pf}"]["cta0_latencies"] = {k: round(v, 1) for k, v in lat.items()}
    if mode == 8 and pf in (0, 96):
        base = ev[0][1000]
        out[f"timeline_mode8_pf{pf}"] = [{"g": int(gg), "mma_pass_empty": int(ev[0][gg] - base), "commit": int(ev[1][gg] - base),
                                  "epi_full_seen": int(ev[2][gg] - base), "epi_release": int(ev[3][gg] - base),
                                  "epi_done": int(ev[4][gg] - base)} for gg in range(1000, 1016)]
ctx.set_knn_mode(0)
ctx.set_knn_probe_flags(0)
print(json.dumps(out, indent=1))
Path("gpurun_out").mkdir(exist_ok=True)
Path("gpurun_out/r02_prof_ts.json").write_text(json.dumps(out, indent=1))
