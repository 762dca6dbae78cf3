"""Timing of the 2-D radius matcher on a reconstruct-sized batch, with the
reference's numpy expression on the host cores beside it.  Prints one JSON line.

    python tests/bench_radius.py [n_seg tracks_per_seg pts_per_seg]

Workload: adding one image to a reconstruction = one segment per already-constructed image that
shares a pair with it (35 for the 36-image configs[1]); each segment tests all tracks that contain
the other image against that pair's verified correspondences.
"""
import json
import sys
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from sfm_b200 import lib as L          # noqa: E402
from oracle import tracks as otr       # noqa: E402  (cpu_baseline leg only)


def main():
    n_seg, nt, npnt = (int(a) for a in sys.argv[1:4]) if len(sys.argv) >= 4 else (35, 20000, 8192)
    rng = np.random.default_rng(3)
    ctx = L.Context(0)
    tracks = rng.uniform(0, [1024, 768], (n_seg * nt, 2)).astype(np.float32).astype(np.float64)
    pts = rng.uniform(0, [1024, 768], (n_seg * npnt, 2)).astype(np.float32).astype(np.float64)
    toff = np.arange(n_seg + 1, dtype=np.int64) * nt
    poff = np.arange(n_seg + 1, dtype=np.int64) * npnt
    d_t, d_p = torch.from_numpy(tracks).cuda(), torch.from_numpy(pts).cuda()
    for _ in range(3):
        so, ti, pi = ctx.radius_match(toff, d_t, poff, d_p, 2.0)
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(10)]
    l0 = ctx.launches
    for e0, e1 in ev:
        e0.record()
        ctx.radius_match(toff, d_t, poff, d_p, 2.0)
        e1.record()
    torch.cuda.synchronize()
    launches = (ctx.launches - l0) // len(ev)
    ms = float(np.median([a.elapsed_time(b) for a, b in ev]))
    tests = float(n_seg) * nt * npnt
    # host-buffer call: tracks and points from pinned memory, hit lists back to the host
    h_t, h_p = torch.from_numpy(tracks).pin_memory(), torch.from_numpy(pts).pin_memory()
    t0 = time.perf_counter()
    for _ in range(5):
        a, b = h_t.cuda(non_blocking=True), h_p.cuda(non_blocking=True)
        so, ti, pi = ctx.radius_match(toff, a, poff, b, 2.0)
        ti.cpu(), pi.cpu()
    e2e_ms = (time.perf_counter() - t0) / 5 * 1e3
    # the reference's expression on the host cores, bounded sample: 1 segment, 2000 tracks
    samp = min(nt, 2000)
    t0 = time.perf_counter()
    reps = 0
    while time.perf_counter() - t0 < 10.0:
        otr.radius_match_numpy(tracks[:samp], pts[:npnt].astype(np.float32), 2.0)
        reps += 1
    cpu_s = (time.perf_counter() - t0) / reps
    cpu_rate = samp * npnt / cpu_s
    # the counting pass evaluates every (track, point) once: 3 f64 instructions (sub, mul, compare;
    # 4 more for the ~0.4 % inside the x band), none of them an FMA; the fill pass copies kept hits.
    # The FP64 pipe issues one instruction per FMA slot: instruction peak = nominal 37 TFLOP/s / 2.
    flops = 3 * tests / (ms * 1e-3)
    peak = ctx.probe_fp64_rate(8192) / 1e12      # measured: all-SM DFMA loop, T f64 instructions/s
    print(json.dumps({
        "metric": "track_point_tests_per_sec", "value": tests / (ms * 1e-3), "unit": "tests/s", "ms": ms,
        "workload": f"{n_seg} segments x {nt} tracks x {npnt} points, threshold 2.0", "hits": int(so[-1]),
        "gpu_launches_per_call": int(launches),
        "e2e": {"value": tests / (e2e_ms * 1e-3), "unit": "tests/s", "ms": e2e_ms,
                "h2d_bytes": int(tracks.nbytes + pts.nbytes), "d2h_bytes": int(8 * so[-1] + so.nbytes)},
        "roofline": {"bound": "fp64 pipe", "achieved": flops / 1e12, "peak": peak, "unit": "T f64 instr/s",
                     "frac": flops / 1e12 / peak, "frac_of_nominal_18.5": flops / 1e12 / 18.5,
                     "peak_source": "measured in this run: sfm_probe_fp64_rate (eight DFMA chains per thread on every SM)",
                     "note": "sub, mul, compare per test, the rest only inside the x band (no FMA: each step rounds as numpy "
                             "does); the ms include the scan, the offset read-back and the fill pass; "
                             "peak = nominal B200 FP64 37 TFLOP/s / 2 flops per FMA slot"},
        "cpu_baseline": {"value": cpu_rate, "unit": "tests/s", "cores": 1, "kind": "reference",
                         "sample": f"np.linalg.norm(t[:, None] - p, axis=2) < 2.0 + np.where on {samp} tracks x {npnt} "
                                   "points (utils/sfm_reconstruction.py:211-213), repeated for 10 s"},
    }))


if __name__ == "__main__":
    main()
