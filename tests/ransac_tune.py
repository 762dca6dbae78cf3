"""Diagnostic (not a pytest module): RANSAC parity statistics on the 148 golden
pairs for a sweep of parameters.  Run on the GPU box."""
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle import fundamental as fm  # noqa: E402
from sfm_b200 import lib as L  # noqa: E402


def main():
    z = np.load(ROOT / "tests/golden/bunny_fundamental.npz")
    off = z["offsets"]
    ctx = L.Context(0)
    p1 = torch.from_numpy(z["pts1"]).cuda()
    p2 = torch.from_numpy(z["pts2"]).cuda()
    P = len(off) - 1
    import itertools
    for thr, lo, mode, seed, mi in itertools.product((3.0,), (2, 4, 8), (3,), (0x5F3759DF2024, 1, 2), (1 << 18, 1 << 16, 1 << 14, 0)):
        if True:
            prm = L.default_params(ransac_thresh=thr, lo_iters=lo, score_mode=mode, seed=seed, min_work=mi)
            o = ctx.ransac_points(off, p1, p2, prm)
            torch.cuda.synchronize()
            mask = o["mask"].cpu().numpy().astype(bool)
            status = o["status"].cpu().numpy()
            met = L.metrics_to_numpy(o["metrics"])
            jac, dI, bad1, badj, flips = [], [], [], [], []
            for k in range(P):
                s = slice(int(off[k]), int(off[k + 1]))
                M = s.stop - s.start
                m_ref = z["mask"][s]
                I_new, I_ref = int(mask[s].sum()), int(m_ref.sum())
                j = (mask[s] & m_ref).sum() / max(1, (mask[s] | m_ref).sum())
                jac.append(j)
                dI.append(I_new - I_ref)
                if I_new < I_ref - max(3, 0.10 * M):
                    bad1.append((str(z["names"][k]), I_new, I_ref, M))
                if M >= 60 and j < 0.85:
                    badj.append((str(z["names"][k]), round(float(j), 3), M))
                if (status[k] == L.PAIR_OK) != bool(z["passed"][k]):
                    flips.append((str(z["names"][k]), I_new, I_ref, M, int(status[k]),
                                  round(float(met["reprojection_error"][k]), 3), int(met["well_distributed"][k]),
                                  round(float(min(z["pts1"][s][mask[s]].std(0).min(), z["pts2"][s][mask[s]].std(0).min())), 2)))
            jac = np.array(jac)
            print(f"lo {lo} seed {seed & 0xff} mi {mi}: jac med {np.median(jac):.3f} p10 {np.percentile(jac, 10):.3f} min {jac.min():.3f} | "
                  f"dI med {np.median(dI):+.0f} min {min(dI)} max {max(dI)} | iters med {np.median(met['ransac_iters']):.0f} | "
                  f"bound-fails {len(bad1)} jac-fails {len(badj)} flips {len(flips)}")
            if bad1 or badj:
                print(f"    bound fails {bad1}; jac fails {badj}")


if __name__ == "__main__":
    main()
