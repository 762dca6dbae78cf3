"""Large BASELINE configs on one GPU (B200 box; not a pytest module): device-resident pass of the
whole hot path over every pair, chunked by the engine.  argv: n_images [n_desc]"""
import json
import sys
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sfm_b200 import lib as L, synth  # noqa: E402
from sfm_b200.engine import PairEngine  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 200
nd = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
s = synth.make_image_set(n, nd, seed=synth.BASE_SEED + 2, device="cuda", tie_frac=0.01)
eng = PairEngine(0)
eng.set_images(s.desc, s.kps, s.offsets.numpy())
pairs = np.array([(i, j) for i in range(n) for j in range(i + 1, n)], np.int32)
torch.cuda.synchronize()
out = {"images": n, "descriptors": nd, "pairs": len(pairs)}
for rep in range(2):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    n_ok = 0
    for lo, hi in eng._chunks(pairs):
        d = eng._match_chunk(pairs[lo:hi])
        n_ok += int((d["status"] == L.PAIR_OK).sum().item())
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    out[f"pass{rep}"] = {"ms": round(ms, 1), "pairs_per_s": round(len(pairs) / (ms * 1e-3), 1),
                         "int8_tops_whole_step": round(256.0 * nd * nd * len(pairs) / (ms * 1e-3) / 1e12, 1),
                         "wall_s": round(time.perf_counter() - t0, 2), "pairs_ok": n_ok}
out["brute_rows"] = eng.ctx.brute_rows()
out["peak_mem_gb"] = round(torch.cuda.max_memory_allocated() / 2**30, 2)
print(json.dumps(out))
