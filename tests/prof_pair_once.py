"""One timed-size launch of a kNN mode for ncu (B200 box; not a pytest module): argv = mode [n_img]."""
import sys
from pathlib import Path
import numpy as np, torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sfm_b200 import lib as L, synth  # noqa: E402
mode = int(sys.argv[1]); n_img = int(sys.argv[2]) if len(sys.argv) > 2 else 13
s = synth.make_image_set(n_img, 8192, seed=synth.BASE_SEED + 1, device="cuda", tie_frac=0.01)
ctx = L.Context(0)
ctx.set_images(s.desc, s.kps, s.offsets.numpy())
pairs = np.array([(i, j) for i in range(n_img) for j in range(i + 1, n_img)], np.int32)
ro = ctx.query_rows(pairs)
buf = ctx.empty((int(ro[-1]), 4), torch.int32)
ctx.set_knn_mode(mode)
for _ in range(2):
    ctx.knn2(pairs, ro, buf)
torch.cuda.synchronize()
print("done", mode, len(pairs))
