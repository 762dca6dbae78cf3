"""GPU bring-up diagnostics (run on the B200 box, not a pytest module).

Checks the TMA + tcgen05 path in isolation (one raw dot-product tile), then the
fused kNN kernel on a small set, and measures the MMA-only int8 rate.  The
checker is a torch fp32 matmul (exact for these integer ranges) -- diagnostics
only; parity proper lives in tests/test_gpu_*.py against the oracle.
"""
import json
import sys
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sfm_b200 import lib as L  # noqa: E402


def ref_knn2(d1, d2):
    a, b = d1.float(), d2.float()
    D = (a * a).sum(1)[:, None] + (b * b).sum(1)[None, :] - 2.0 * (a @ b.T)
    D = D.round().to(torch.int64)
    n = D.shape[1]
    key = D * (1 << 20) + torch.arange(n, device=D.device)[None, :]
    k2, _ = torch.topk(key, 2, dim=1, largest=False)
    return (k2 & ((1 << 20) - 1)).int(), (k2 >> 20).int()


def main():
    torch.backends.cuda.matmul.allow_tf32 = False
    out = {}
    ctx = L.Context(0)
    g = torch.Generator(device="cuda").manual_seed(1)
    n_img, n = 3, 1000
    desc = torch.randint(0, 256, (n_img * n, 128), dtype=torch.uint8, device="cuda", generator=g)
    kps = torch.rand(n_img * n, 2, device="cuda", generator=g) * 1000
    off = np.arange(n_img + 1, dtype=np.int64) * n
    ctx.set_images(desc, kps, off)
    torch.cuda.synchronize()

    # 1. raw tile
    for (q, t) in [(0, 0), (5, 1000), (131, 77)]:
        got = ctx.debug_dot_tile(q, t)
        torch.cuda.synchronize()
        want = (desc[q:q + 128].float() @ desc[t:t + 128].float().T).round().int()
        bad = (got != want)
        out[f"dot_tile_{q}_{t}_mismatch"] = int(bad.sum())
        if bad.any():
            idx = bad.nonzero()[:8].tolist()
            print("dot tile mismatch", q, t, "count", int(bad.sum()), "first", idx)
            print("got", got[:4, :8].tolist())
            print("want", want[:4, :8].tolist())
            # pattern help: compare against transposes / row permutations
            print("rows all-bad?", bad.all(1).sum().item(), "cols all-bad?", bad.all(0).sum().item())
    print(json.dumps(out))

    # 2. fused kNN
    pairs = np.array([(0, 1), (1, 2), (2, 0), (1, 0)], np.int32)
    knn, ro = ctx.knn2(pairs)
    torch.cuda.synchronize()
    for p, (qi, ti) in enumerate(pairs):
        idx, d2 = ref_knn2(desc[off[qi]:off[qi + 1]], desc[off[ti]:off[ti + 1]])
        got = knn[ro[p]:ro[p + 1]]
        out[f"knn_pair{p}_idx_mismatch"] = int((got[:, :2] != idx).sum())
        out[f"knn_pair{p}_d2_mismatch"] = int((got[:, 2:] != d2).sum())
    print(json.dumps(out))

    # 3. int8 MMA-only rate
    for iters in (2048, 8192):
        tops, ms = ctx.probe_int8_peak(iters)
        out[f"int8_probe_tops_{iters}"] = round(tops, 1)
        out[f"int8_probe_ms_{iters}"] = round(ms, 3)
    print(json.dumps(out))

    # 4. kNN throughput, 36 x 8192
    n_img, n = 36, 8192
    desc = torch.randint(0, 256, (n_img * n, 128), dtype=torch.uint8, device="cuda", generator=g)
    off = np.arange(n_img + 1, dtype=np.int64) * n
    ctx.set_images(desc, None, off)
    pairs = np.array([(i, j) for i in range(n_img) for j in range(i + 1, n_img)], np.int32)
    ro = ctx.query_rows(pairs)
    buf = ctx.empty((int(ro[-1]), 4), torch.int32)
    for _ in range(2):
        ctx.knn2(pairs, ro, buf)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        ctx.knn2(pairs, ro, buf)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    ops = 256.0 * n * n * len(pairs)
    out["knn_36x8k_ms"] = round(ms, 3)
    out["knn_36x8k_tops"] = round(ops / ms / 1e9, 1)
    out["knn_36x8k_pairs_per_s"] = round(len(pairs) / ms * 1e3, 1)
    idx, d2 = ref_knn2(desc[:n], desc[n:2 * n])
    got = buf[:n]
    out["knn_8k_idx_mismatch"] = int((got[:, :2] != idx).sum())
    out["knn_8k_d2_mismatch"] = int((got[:, 2:] != d2).sum())
    print(json.dumps(out))
    Path("gpurun_out").mkdir(exist_ok=True)
    Path("gpurun_out/diag.json").write_text(json.dumps(out, indent=1))


if __name__ == "__main__":
    t = time.time()
    main()
    print("diag done in %.1fs" % (time.time() - t))
