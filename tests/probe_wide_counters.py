"""Where the wide kNN kernel's warps wait (B200 box; not a pytest module): probe bit 4 makes CTA 0 count cycles
in its MMA warp and its first epilogue warp."""
import json
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sfm_b200 import lib as L, synth  # noqa: E402

n_img, n_desc = 24, 8192
s = synth.make_image_set(n_img, n_desc, seed=synth.BASE_SEED + 1, device="cuda", tie_frac=0.01)
ctx = L.Context(0)
ctx.set_images(s.desc, s.kps, s.offsets.numpy())
pairs = np.array([(i, j) for i in range(n_img) for j in range(i + 1, n_img)], np.int32)
ro = ctx.query_rows(pairs)
buf = ctx.empty((int(ro[-1]), 4), torch.int32)
out = {}
pf_extra = int(sys.argv[2]) if len(sys.argv) > 2 else 0
for mode in [int(m) for m in (sys.argv[1].split(",") if len(sys.argv) > 1 else "5,6,7,17,18,19".split(","))]:
    ctx.set_knn_mode(mode)
    ctx.set_knn_probe_flags(16 | pf_extra)
    ctx.knn2(pairs, ro, buf)
    torch.cuda.synchronize()
    c = ctx.pair_counters()
    tiles = max(c[7], 1)
    out[f"mode{mode}"] = {
        "epi_loop_cycles": c[0], "epi_wait_full": c[1], "epi_wait_tmem_ld": c[2], "epi_end_of_item": c[3],
        "mma_loop_cycles": c[4], "mma_wait_b_full": c[5], "mma_wait_t_empty": c[6], "tiles": c[7],
        "per_tile": {"epi_total": round(c[0] / tiles, 1), "epi_wait_full": round(c[1] / tiles, 1),
                     "epi_wait_tmem_ld": round(c[2] / tiles, 1), "epi_end_of_item": round(c[3] / tiles, 1),
                     "mma_total": round(c[4] / tiles, 1), "mma_wait_b_full": round(c[5] / tiles, 1),
                     "mma_wait_t_empty": round(c[6] / tiles, 1)}}
    print(mode, json.dumps(out[f"mode{mode}"]["per_tile"]), flush=True)
ctx.set_knn_mode(0)
ctx.set_knn_probe_flags(0)
Path("gpurun_out").mkdir(exist_ok=True)
Path(f"gpurun_out/r02_probe_wide_counters_pf{pf_extra}.json").write_text(json.dumps(out, indent=1))
