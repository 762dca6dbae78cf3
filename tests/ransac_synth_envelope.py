"""Diagnostic (not a pytest module; B200 box): the SURVEY 8c RANSAC envelope on SYNTHETIC 8k-config pairs
(the bench workload's shape: 8192 descriptors per image, median ~450 matches per pair) against
cv2.findFundamentalMat run live on the same correspondences (the reference's call,
utils/find_matches.py:282), for a sweep of the hypothesis floor `min_work`, with the RANSAC kernel's time."""
import json
import sys
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle import fundamental as fm  # noqa: E402
from sfm_b200 import lib as L, synth  # noqa: E402
from sfm_b200.engine import PairEngine  # noqa: E402


def envelope(off, pts1, pts2, mask, status, ref, passing_only=True):
    jac, dI, bound, jfail, flips = [], [], 0, 0, 0
    for k in range(len(off) - 1):
        if ref[k] is None:
            continue
        if passing_only and not ref[k]["passed"]:
            flips += status[k] == L.PAIR_OK
            continue
        s = slice(int(off[k]), int(off[k + 1]))
        M = s.stop - s.start
        m_ref = ref[k]["mask"]
        I_new, I_ref = int(mask[s].sum()), int(m_ref.sum())
        u = (mask[s] | m_ref).sum()
        j = (mask[s] & m_ref).sum() / u if u else 1.0
        jac.append(j)
        dI.append(I_new - I_ref)
        bound += I_new < I_ref - max(3, 0.10 * M)
        jfail += M >= 60 and j < 0.85
        flips += (status[k] == L.PAIR_OK) != ref[k]["passed"]
    jac = np.array(jac)
    return {"pairs": len(jac), "jaccard_median": round(float(np.median(jac)), 4), "jaccard_p10": round(float(np.percentile(jac, 10)), 4),
            "jaccard_min": round(float(jac.min()), 4), "dI_median": float(np.median(dI)), "dI_min": int(min(dI)),
            "bound_fails": int(bound), "jaccard_fails": int(jfail), "gate_flips": int(flips)}


def main():
    n_img = int(sys.argv[1]) if len(sys.argv) > 1 else 21
    s = synth.make_image_set(n_img, 8192, seed=synth.BASE_SEED + 2, tie_frac=0.01)
    eng = PairEngine(0)
    eng.upload_images(s.desc, s.kps, s.offsets.numpy())
    pairs = synth.all_pairs(n_img).numpy()
    res = eng.match_pairs(pairs, only_ok=False)
    off = res.pack_off
    t0 = time.time()
    ref = []
    for k in range(len(pairs)):
        sl = res.pair_slice(k)
        p1, p2 = res.pts1[sl], res.pts2[sl]
        if len(p1) < 8:
            ref.append(None)
            continue
        F, _ = fm.find_fundamental_cv2(p1, p2)
        if F is None or np.asarray(F).shape != (3, 3):
            ref.append(None)
            continue
        v = fm.geometric_verification(p1, p2, F)
        ref.append({"mask": v["inlier_mask"], "passed": bool(fm.verify_match_quality(v))})
    cv_s = time.time() - t0
    out = {"n_images": n_img, "pairs": len(pairs), "matches_median": float(np.median(res.count)),
           "cv2_seconds": round(cv_s, 2), "ref_passing": sum(1 for r in ref if r and r["passed"])}
    d1 = torch.from_numpy(res.pts1).cuda()
    d2 = torch.from_numpy(res.pts2).cuda()
    # timing on the same correspondences repeated 24 times (5040 pairs: eleven waves of 444 resident CTAs)
    reps = 24
    cnt = np.diff(off)
    off_rep = np.concatenate([[0], np.cumsum(np.tile(cnt, reps))]).astype(np.int64)
    r1, r2 = d1.repeat(reps, 1), d2.repeat(reps, 1)
    sweep = [(mw, 3) for mw in (1 << 18, 0)] + [(1 << 18, sm) for sm in (2, 1, 0)]
    for mw, sm in sweep:
        for seed in (0x5F3759DF2024, 7):
            prm = L.default_params(min_work=mw, seed=seed, score_mode=sm)
            eng.ctx.ransac_points(off_rep, r1, r2, prm)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            eng.ctx.ransac_points(off_rep, r1, r2, prm)
            e1.record()
            torch.cuda.synchronize()
            o = eng.ctx.ransac_points(off, d1, d2, prm)
            torch.cuda.synchronize()
            env = envelope(off, res.pts1, res.pts2, o["mask"].cpu().numpy().astype(bool), o["status"].cpu().numpy(), ref)
            met = L.metrics_to_numpy(o["metrics"])
            env["us_per_pair"] = round(1e3 * e0.elapsed_time(e1) / (reps * len(pairs)), 3)
            env["iters_median"] = float(np.median(met["ransac_iters"]))
            env["iters_mean"] = float(np.mean(met["ransac_iters"]))
            out[f"min_work_{mw}_mode_{sm}_seed_{seed & 0xff}"] = env
            print(json.dumps({f"min_work_{mw}_mode_{sm}_seed_{seed & 0xff}": env}), flush=True)
    Path("gpurun_out").mkdir(exist_ok=True)
    Path("gpurun_out/r02_ransac_synth_envelope.json").write_text(json.dumps(out, indent=1))
    print(json.dumps({k: v for k, v in out.items() if not k.startswith("min_work")}))


if __name__ == "__main__":
    main()
