#!/usr/bin/env python
"""Benchmark of the pairwise-matching hot path (BASELINE.json metric: image
pairs/s and descriptor comparisons/s as int8 TOPS vs peak).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

A step is one pass of the hot path (kNN k=2 -> ratio -> RANSAC F -> verification
-> result packing) over every pair of a synthetic SIFT-like image set:
  N = 1   36 images x 8192 descriptors, 630 pairs   (BASELINE configs[1])
  N > 1   200 images x 8192 descriptors, 19,900 pairs partitioned over the ranks
          (BASELINE configs[2]); descriptors replicated with one NCCL all-gather.
`value` is device-resident throughput (CUDA events, max over ranks); `e2e` goes
through the public engine call with host buffers (H2D + D2H inside the timed
region).  `--impl reference` times the reference's own OpenCV CPU path (via the
oracle, which restates the reference's cv2 calls) on the host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

N_DESC = 8192
NOMINAL_INT8_TOPS = 4500.0          # B200 dense int8 datasheet figure
OPS_PER_CMP = 256.0                 # 2 * 128 int8 ops per descriptor comparison


def workload(n_gpus: int):
    return dict(name="synthetic 200 images x 8192 SIFT descriptors, exhaustive 19900 pairs", n_images=200)


def all_pairs(n):
    return np.array([(i, j) for i in range(n) for j in range(i + 1, n)], np.int32)


# --------------------------------------------------------------------- clocks
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows = []
        self.proc = None
        self.idx = gpu_index

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20",
                 "-i", str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def mark(self):
        """Samples before this point (sampler start-up, warm-up) are not part of the summary."""
        self.rows = self.rows[len(self.rows):]
        self.t_mark = time.perf_counter()

    def extend_until(self, n_samples: int, load, limit_s: float = 6.0):
        """nvidia-smi needs a moment to start (longer with 8 pollers on one box): if the timed region
        was shorter than that, keep the same load running, untimed, until n_samples rows exist."""
        t0 = time.perf_counter()
        extended = False
        while self.proc and len(self.rows) < n_samples and time.perf_counter() - t0 < limit_s:
            load()
            extended = True
        return extended

    def __exit__(self, *a):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
                for k, nme in enumerate(names):
                    if r[4 + k].lower().startswith("active"):
                        reasons.add(nme)
            except Exception:
                continue
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        load = sorted(sm)[len(sm) // 2:]
        return {"sm_mhz": float(np.median(load)), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------- CPU reference arm
def cpu_reference_sample(desc, kps, offsets, pairs, budget_s: float, max_pairs: int):
    """The reference's OpenCV path (BFMatcher.knnMatch + ratio loop + findFundamentalMat +
    geometric_verification + verify_match_quality; utils/find_matches.py:141-214,
    :272-290) on a bounded sample of the workload's pairs.  Returns (pairs_done, seconds, cores)."""
    import cv2
    from oracle import pipeline as opipe
    cores = os.cpu_count() or 1
    cv2.setNumThreads(cores)
    rng = np.random.default_rng(20240)
    order = rng.permutation(len(pairs))[:max_pairs]
    t0 = time.perf_counter()
    done = 0
    for k in order:
        i, j = pairs[k]
        a, b = slice(offsets[i], offsets[i + 1]), slice(offsets[j], offsets[j + 1])
        opipe.process_pair(desc[a], kps[a], desc[b], kps[b], mutual=False, use_cv2=True)
        done += 1
        if time.perf_counter() - t0 > budget_s:
            break
    return done, time.perf_counter() - t0, min(cores, cv2.getNumThreads())


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    from sfm_b200 import synth
    wl = workload(args.gpus)
    n = wl["n_images"]
    # the same synthetic set as the B200 arm (seeded), generated on the CPU for a sample of images
    n_gen = min(n, 12)
    s = synth.make_image_set(n_gen, N_DESC, seed=synth.BASE_SEED + (1 if n == 36 else 2), tie_frac=0.01)
    desc, kps, off = s.desc.numpy(), s.kps.numpy(), s.offsets.numpy()
    pairs = all_pairs(n_gen)
    per_step = max(2, min(8, len(pairs) // max(1, args.steps + args.warmup)))
    rng = np.random.default_rng(7)
    order = rng.permutation(len(pairs))
    import cv2
    from oracle import pipeline as opipe
    cores = os.cpu_count() or 1
    cv2.setNumThreads(cores)

    def step(k):
        for idx in order[(k * per_step) % len(order):][:per_step]:
            i, j = pairs[idx]
            a, b = slice(off[i], off[i + 1]), slice(off[j], off[j + 1])
            opipe.process_pair(desc[a], kps[a], desc[b], kps[b], mutual=False, use_cv2=True)

    for k in range(args.warmup):
        step(k)
    t0 = time.perf_counter()
    for k in range(args.steps):
        step(args.warmup + k)
    dt = time.perf_counter() - t0
    pairs_s = per_step * args.steps / dt
    cmp_s = pairs_s * N_DESC * N_DESC
    line = {
        "impl": "reference", "metric": "image_pairs_per_sec", "value": pairs_s, "unit": "pairs/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl["name"], "mutual_nn": False, "ratio": 0.75,
                   "note": f"each step = {per_step} pairs sampled from the workload (bounded CPU sample)"},
        "comparisons_per_sec": cmp_s, "int8_equiv_tops": cmp_s * OPS_PER_CMP / 1e12,
        "cpu_baseline": {"value": pairs_s, "unit": "pairs/s", "cores": min(cores, cv2.getNumThreads()),
                         "kind": "port",
                         "sample": f"{per_step * args.steps} pairs of 8192x8192 descriptors through the reference's own "
                                   f"OpenCV calls (cv2 {cv2.__version__} BFMatcher.knnMatch + ratio loop + findFundamentalMat "
                                   "+ geometric_verification), restated by oracle/pipeline.py"},
        "e2e": {"value": pairs_s, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_b200(args):
    import torch
    import torch.distributed as dist
    from sfm_b200 import dist as D
    from sfm_b200 import lib as L
    from sfm_b200 import synth
    from sfm_b200.engine import PairEngine

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback for the B200 arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    wl = workload(world)
    n = wl["n_images"]
    seed = synth.BASE_SEED + (1 if n == 36 else 2)
    s = synth.make_image_set(n, N_DESC, seed=seed, device=dev, tie_frac=0.01)   # synthetic, untimed
    off = s.offsets.numpy()
    counts = np.diff(off)
    pairs = all_pairs(n)
    lo, hi = D.partition_pairs(pairs, counts, world)[rank]
    my_pairs = pairs[lo:hi]
    eng = PairEngine(local_rank, mutual=bool(args.mutual))
    eng.set_images(s.desc, s.kps, off)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)               # > 126 MB L2

    # ---------------------------------------------------------- device-resident steps
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    with ClockSampler(local_rank) as clk:               # started before the warm-up so that it is polling by now
        for _ in range(args.warmup):
            eng.match_device(my_pairs)
        launches0 = eng.ctx.launches
        barrier()
        clk.mark()
        for e0, e1 in ev:
            flush.zero_()
            e0.record()
            chunks = eng.match_device(my_pairs)
            e1.record()
        barrier()
    launches = eng.ctx.launches - launches0
    ms_total = max_over_ranks(sum(e0.elapsed_time(e1) for e0, e1 in ev))
    ms_step = ms_total / args.steps
    n_ok = sum(int((c["status"] == L.PAIR_OK).sum().item()) for c in chunks)
    if world > 1:
        t_ok = torch.tensor([n_ok], dtype=torch.int64, device=dev)
        dist.all_reduce(t_ok)
        n_ok = int(t_ok.item())
    pairs_s = len(pairs) / (ms_step * 1e-3)
    cmp_s = float((counts[pairs[:, 0]].astype(np.float64) * counts[pairs[:, 1]]).sum()) / (ms_step * 1e-3)

    # ---------------------------------------------------------- dominant kernel alone (roofline)
    clk_rows_step = list(clk.rows)
    first = next(iter(eng._chunks(my_pairs)))
    kp = my_pairs[first[0]:first[1]]
    ro = eng.ctx.query_rows(kp)
    kbuf = eng._buf("knn", (int(ro[-1]), 4), torch.int32)
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    with ClockSampler(local_rank) as clk_k:
        for e0, e1 in kev:
            flush.zero_()
            e0.record()
            eng.ctx.knn2(kp, ro, kbuf)
            e1.record()
        torch.cuda.synchronize()

        def same_load():
            eng.ctx.knn2(kp, ro, kbuf)
            torch.cuda.synchronize()
        clk_k.rows = clk_rows_step + clk_k.rows    # clocks under load: timed steps + kernel-alone region
        clk_extended = clk_k.extend_until(5, same_load)
    clk.rows = clk_k.rows
    k_ms = float(np.mean([e0.elapsed_time(e1) for e0, e1 in kev]))
    k_ops = OPS_PER_CMP * float((counts[kp[:, 0]].astype(np.float64) * counts[kp[:, 1]]).sum())
    k_tops = k_ops / (k_ms * 1e-3) / 1e12
    probe_tops, _ = eng.ctx.probe_int8_peak(4096)

    # ---------------------------------------------------------- end to end through the public call
    blocks = D.image_blocks(n, world)
    b0, b1 = blocks[rank]
    h_desc = s.desc[off[b0]:off[b1]].cpu().pin_memory()
    h_kps = s.kps[off[b0]:off[b1]].cpu().pin_memory()
    my_counts = counts[b0:b1]
    h2d = h_desc.numel() + h_kps.numel() * 4

    def e2e_step():
        d = h_desc.to(dev, non_blocking=True)
        k = h_kps.to(dev, non_blocking=True)
        if world > 1:
            d, k, o = D.replicate_images(d, k, my_counts)      # NCCL all-gather over NVLink
        else:
            o = off
        eng.set_images(d, k, o)
        if world == 1:
            res = eng.match_pairs(my_pairs)                    # host-side results (D2H inside)
            return sum(getattr(res, f).nbytes for f in ("status", "count", "metrics", "F", "pack_off", "query_idx",
                                                        "train_idx", "sq_dist", "inlier_mask", "sym_err", "pts1",
                                                        "pts2"))
        # N > 1, as ImageMatcher's rank path: every rank brings the match lists of ITS pairs to its own
        # host memory (it writes those files); only the per-pair status / metrics rows meet on rank 0
        res = eng.match_pairs(my_pairs)
        nbytes = sum(getattr(res, f).nbytes for f in ("status", "count", "metrics", "F", "pack_off", "query_idx",
                                                      "train_idx", "sq_dist", "inlier_mask", "sym_err", "pts1", "pts2"))
        rows = torch.from_numpy(np.ascontiguousarray(res.metrics).view(np.uint8).reshape(len(my_pairs), -1)).to(dev)
        D.gather_to_root(rows)
        D.gather_to_root(torch.from_numpy(np.ascontiguousarray(res.status)).to(dev))
        return nbytes

    e2e_step()
    barrier()
    t0 = time.perf_counter()
    d2h = 0
    for _ in range(args.steps):
        d2h = e2e_step()
    barrier()
    e2e_s = max_over_ranks(time.perf_counter() - t0) / args.steps
    if world > 1:                                              # bytes of all ranks
        t_b = torch.tensor([d2h, h2d], dtype=torch.int64, device=dev)
        dist.all_reduce(t_b)
        d2h, h2d = int(t_b[0].item()), int(t_b[1].item())
    e2e_pairs_s = len(pairs) / e2e_s

    # ---------------------------------------------------------- CPU baseline (rank 0, N = 1)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        nd = 6
        done, secs, cores = cpu_reference_sample(s.desc[:off[nd]].cpu().numpy(), s.kps[:off[nd]].cpu().numpy(), off,
                                                 all_pairs(nd), budget_s=20.0, max_pairs=15)
        cpu = {"value": done / secs, "unit": "pairs/s", "cores": cores, "kind": "port",
               "sample": f"{done} pairs of the workload (first {nd} images, 8192x8192 descriptors each) through the "
                         "reference's OpenCV calls restated in oracle/pipeline.py (BFMatcher.knnMatch k=2 + ratio loop + "
                         "findFundamentalMat FM_RANSAC + geometric_verification)"}

    if rank == 0:
        peaks = {}
        try:
            peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
        except Exception:
            pass
        traffic = None
        tp = ROOT / "profiles" / "knn2_traffic.json"
        if tp.exists():
            try:
                traffic = json.loads(tp.read_text()).get("dram_bytes_per_launch")
            except Exception:
                traffic = None
        line = {
            "metric": "image_pairs_per_sec", "value": pairs_s, "unit": "pairs/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": wl["name"], "mutual_nn": bool(args.mutual), "ratio": 0.75,
                       "pairs": int(len(pairs)), "pairs_passing_verification": n_ok,
                       "parallelism": f"pair-partition x{world}", "l2": "flushed between steps (256 MiB write)",
                       "note": "N=1 runs BASELINE configs[1] (36x8k); N>1 runs configs[2] (200x8k) split over the ranks"},
            "comparisons_per_sec": cmp_s, "int8_tops": cmp_s * OPS_PER_CMP / 1e12,
            "roofline": {"bound": "tensor", "achieved": k_tops, "peak": NOMINAL_INT8_TOPS, "unit": "TOP/s",
                         "frac": k_tops / NOMINAL_INT8_TOPS, "traffic": traffic,
                         "kernel": "knn2_wide_kernel (TMA + tcgen05 kind::i8 128x256x32, fused top-2 epilogue) + its fix-up pass",
                         "kernel_ms": k_ms,
                         "peak_source": "nominal dense int8 (MEASURED_PEAKS.json has no int8 entry); "
                                        f"MMA-only tcgen05 probe measured live: {probe_tops:.0f} TOP/s; "
                                        f"MEASURED bf16 {peaks.get('bf16_tflops')} TFLOP/s, HBM {peaks.get('hbm_gbs')} GB/s",
                         "probe_tops": probe_tops},
            "e2e": {"value": e2e_pairs_s, "unit": "pairs/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)},
            "gpu_launches": int(launches),
            "clocks": dict(clk.summary(), window="timed steps + kernel-alone region" +
                           (" + untimed continuation of the same kernel (sampler start-up)" if clk_extended else "")),
        }
        if cpu:
            line["cpu_baseline"] = cpu
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--mutual", type=int, default=0, help="1: add the mutual-NN check (second, reversed kNN pass)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.steps = max(1, args.steps)
    if args.impl != "reference":
        args.warmup = max(3, args.warmup)     # timing rule: at least three warm-up steps; the line reports what ran
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
