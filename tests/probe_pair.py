"""CTA-pair kNN kernel bring-up (B200 box; not a pytest module): the folded accumulators of a
256 x 128 tile through tcgen05.mma.cta_group::2 against numpy, then the kernel (mode 12) against the
wide kernel (mode 5) on a small and an 8k set, then timings of modes 5 / 12 / 13 / 14."""
import json
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sfm_b200 import lib as L, synth  # noqa: E402

stage = sys.argv[1] if len(sys.argv) > 1 else "tile"
out = {}
ctx = L.Context(0)
C = 255 * 31 * 255 + 254
if stage == "tile":
    s = synth.make_image_set(2, [640, 700], seed=5, device="cuda")
    ctx.set_images(s.desc, s.kps, s.offsets.numpy())
    d = s.desc.cpu().numpy().astype(np.int64)
    for (q, t) in ((0, 640), (128, 640 + 256), (384, 0), (37, 640 + 131)):
        got = ctx.debug_pair_tile(q, t).cpu().numpy()
        a, b = d[q:q + 256], d[t:t + 128]
        want = a @ b.T + (C - (b * b).sum(1) // 2)[None, :]
        ok = bool(np.array_equal(got[:len(a), :len(b)], want))
        out[f"tile_q{q}_t{t}_exact"] = ok
        if not ok:
            bad = np.argwhere(got[:len(a), :len(b)] != want)
            out[f"tile_q{q}_t{t}_bad"] = {"n": int(len(bad)), "rows": sorted(set(bad[:, 0].tolist()))[:8],
                                          "cols": sorted(set(bad[:, 1].tolist()))[:8],
                                          "got": got[:2, :4].tolist(), "want": want[:2, :4].tolist()}
elif stage == "small":
    s = synth.make_image_set(5, [600, 700, 650, 130, 1000], seed=11, device="cuda", tie_frac=0.02)
    ctx.set_images(s.desc, s.kps, s.offsets.numpy())
    pairs = np.array([(i, j) for i in range(5) for j in range(5) if i != j], np.int32)
    ro = ctx.query_rows(pairs)
    res = {}
    for mode in (5, 12):
        ctx.set_knn_mode(mode)
        buf = ctx.empty((int(ro[-1]), 4), torch.int32)
        ctx.knn2(pairs, ro, buf)
        torch.cuda.synchronize()
        res[mode] = buf.cpu().numpy()
    out["small_equal"] = bool(np.array_equal(res[5], res[12]))
    if not out["small_equal"]:
        bad = np.argwhere((res[5] != res[12]).any(1))[:, 0]
        out["small_bad_rows"] = bad[:10].tolist()
        out["small_bad_n"] = int(len(bad))
        out["small_sample"] = [res[5][bad[:3]].tolist(), res[12][bad[:3]].tolist()]
else:
    n_img = int(sys.argv[2]) if len(sys.argv) > 2 else 16
    s = synth.make_image_set(n_img, 8192, seed=synth.BASE_SEED + 1, device="cuda", tie_frac=0.01)
    ctx.set_images(s.desc, s.kps, s.offsets.numpy())
    pairs = np.array([(i, j) for i in range(n_img) for j in range(i + 1, n_img)], np.int32)
    ro = ctx.query_rows(pairs)
    buf = ctx.empty((int(ro[-1]), 4), torch.int32)
    ops = 256.0 * 8192 * 8192 * len(pairs)
    ref = None
    modes = [int(m) for m in sys.argv[3].split(",")] if len(sys.argv) > 3 else [5, 12, 13, 14, 5, 12]
    for mode in modes:
        ctx.set_knn_mode(mode)
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
        out.setdefault(f"mode{mode}_tops", []).append(ts)
        if mode in (5, 12):
            got = buf.clone()
            if ref is None:
                ref = got
            out[f"mode{mode}_equals_first"] = bool(torch.equal(got, ref))
            out[f"brute_rows_mode{mode}"] = ctx.brute_rows()
        print(json.dumps(out), flush=True)
ctx.set_knn_mode(0)
print(json.dumps(out), flush=True)
Path("gpurun_out").mkdir(exist_ok=True)
Path(f"gpurun_out/r02_probe_pair_{stage}.json").write_text(json.dumps(out, indent=1))
