"""Diagnostic (B200 box): where the RANSAC stage of the 200 x 8k bench workload spends its hypotheses."""
import json, sys
from pathlib import Path
import numpy as np, torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sfm_b200 import lib as L, synth  # noqa: E402
from sfm_b200.engine import PairEngine  # noqa: E402
n_img = int(sys.argv[1]) if len(sys.argv) > 1 else 200
s = synth.make_image_set(n_img, 8192, seed=synth.BASE_SEED + 2, tie_frac=0.01)
eng = PairEngine(0)
eng.upload_images(s.desc, s.kps, s.offsets.numpy())
pairs = synth.all_pairs(n_img).numpy()
res = eng.match_pairs(pairs, only_ok=True)
m = res.count.astype(np.int64)
it = res.metrics["ransac_iters"].astype(np.int64)
ok = res.status == L.PAIR_OK
out = {"pairs": len(pairs), "passing": int(ok.sum())}
for name, sel in (("passing", ok), ("failing", ~ok)):
    out[name] = {"n": int(sel.sum()), "m_quantiles": np.percentile(m[sel], [0, 10, 50, 90, 100]).tolist() if sel.any() else [],
                 "iters_quantiles": np.percentile(it[sel], [0, 10, 50, 90, 100]).tolist() if sel.any() else [],
                 "sum_iters": int(it[sel].sum()), "sum_iters_x_m": int((it[sel] * m[sel]).sum())}
bins = [0, 5, 8, 15, 32, 64, 128, 256, 384, 512, 1024, 1 << 20]
out["failing_m_hist"] = {f"[{a},{b})": int(((m >= a) & (m < b) & ~ok).sum()) for a, b in zip(bins[:-1], bins[1:])}
out["passing_m_hist"] = {f"[{a},{b})": int(((m >= a) & (m < b) & ok).sum()) for a, b in zip(bins[:-1], bins[1:])}
out["status_hist"] = {L.STATUS_NAMES[k]: int((res.status == k).sum()) for k in range(5)}
out["inliers_of_failing_quantiles"] = np.percentile(res.metrics["ransac_inliers"][~ok], [0, 50, 90, 99, 100]).tolist()
print(json.dumps(out, indent=1))
Path("gpurun_out/r02_diag_ransac_work.json").write_text(json.dumps(out, indent=1))
