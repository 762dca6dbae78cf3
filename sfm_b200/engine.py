"""Batched pair engine: descriptors in, verified matches out.

Sequences the kernels behind the C ABI for a list of image pairs -- the body of
``ImageMatcher.process_image_pair`` from descriptors on (reference
utils/find_matches.py:272-290) for a whole batch at once:

    K1 sfm_knn2            (:144-147)  kNN k=2, forward (and reversed for mutual-NN)
    K2 sfm_filter_matches  (:150-153)  ratio test (+ mutual) -> compact match lists
    K3/K4 sfm_verify_pairs (:278-290)  gather, RANSAC F, verification, quality gate
    sfm_pack_results                    dense arrays for the device->host copy

PyTorch only owns memory and streams here.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

from . import lib as L

RATIO = 0.75           # utils/find_matches.py:152


@dataclass
class PairBatchResult:
    """Host-side result of one batch (numpy).  Per-pair arrays are length P;
    the dense arrays hold only the kept pairs' matches (see ``pack_off``)."""
    pairs: np.ndarray            # int32 [P,2] image ids
    status: np.ndarray           # int32 [P] lib.PAIR_*
    count: np.ndarray            # int32 [P] matches after ratio(+mutual)
    metrics: np.ndarray          # structured [P] (lib.METRICS_DTYPE)
    F: np.ndarray                # float64 [P,3,3]
    pack_off: np.ndarray         # int64 [P+1] offsets into the dense arrays
    query_idx: np.ndarray        # int32 dense
    train_idx: np.ndarray        # int32 dense
    sq_dist: np.ndarray          # int32 dense
    inlier_mask: np.ndarray      # bool dense
    sym_err: np.ndarray          # float32 dense
    pts1: np.ndarray             # float32 dense [.,2]
    pts2: np.ndarray             # float32 dense [.,2]
    timings_ms: dict = field(default_factory=dict)

    def pair_slice(self, p: int) -> slice:
        return slice(int(self.pack_off[p]), int(self.pack_off[p + 1]))

    def distance(self, p: int) -> np.ndarray:
        """``DMatch.distance`` values: float32 sqrt of the exact integer d^2."""
        return np.sqrt(self.sq_dist[self.pair_slice(p)].astype(np.float32)).astype(np.float32)


class PairEngine:
    """One GPU's matcher: register an image set, then match pair batches."""

    def __init__(self, device: int = 0, ratio: float = RATIO, mutual: bool = False,
                 params: L.VerifyParams | None = None, max_rows_per_chunk: int = 48 << 20):
        import torch
        self.torch = torch
        self.ctx = L.Context(device)
        self.device = self.ctx.device
        self.ratio = float(ratio)
        # mutual: False, True / "sparse" (only the surviving matches' train descriptors are matched back:
        # sfm_mutual_filter) or "full" (a second exhaustive kNN pass over the reversed pairs)
        self.mutual = bool(mutual)
        self.mutual_full = mutual == "full"
        self.params = params if params is not None else L.default_params()
        self.max_rows = int(max_rows_per_chunk)
        self.counts = None
        self._bufs = {}
        self._copy_stream = None     # device -> host copies of match_pairs
        self._marks = None           # stage_times() profiling events

    # ------------------------------------------------------------ image set
    def set_images(self, desc, kps, offsets):
        """desc u8 [R,128], kps f32 [R,2] (device tensors), offsets int64 [N+1] host."""
        self.ctx.set_images(desc, kps, offsets)
        self.counts = self.ctx.counts
        self.n_images = len(self.counts)
        self._sparse_ok = None      # fold-range flag of this set, read when the mutual check first needs it

    def upload_images(self, desc_host, kps_host, offsets):
        """Host (ideally pinned) tensors -> device, then register.  Returns bytes copied."""
        torch = self.torch
        d = desc_host.to(self.device, non_blocking=True)
        k = kps_host.to(self.device, non_blocking=True)
        self.set_images(d, k, offsets)
        return desc_host.numel() * desc_host.element_size() + kps_host.numel() * kps_host.element_size()

    def _buf(self, name, shape, dtype):
        torch = self.torch
        n = int(np.prod(shape))
        cur = self._bufs.get(name)
        if cur is None or cur.numel() < n or cur.dtype != dtype:
            cur = torch.empty(max(n, 1), dtype=dtype, device=self.device)
            self._bufs[name] = cur
        return cur[:n].view(*shape)

    # --------------------------------------------------------------- chunks
    def _chunks(self, pairs):
        """(lo, hi) ranges of the pair list whose kNN rows fit the arenas (at least one pair each)."""
        rows = self.counts[pairs[:, 0]].astype(np.int64)
        if self.mutual and self._full_reverse():
            rows = np.maximum(rows, self.counts[pairs[:, 1]].astype(np.int64))
        cum = np.cumsum(rows)
        lo, n = 0, len(pairs)
        while lo < n:
            before = int(cum[lo - 1]) if lo else 0
            hi = int(np.searchsorted(cum, before + self.max_rows, side="right"))
            hi = min(max(hi, lo + 1), n)
            yield lo, hi
            lo = hi

    def _full_reverse(self) -> bool:
        if self.mutual_full:
            return True
        if self._sparse_ok is None:
            self._sparse_ok = int(self.counts.sum()) == 0 or not self.ctx.fold_flag()
        return not self._sparse_ok

    # ----------------------------------------------------------- device path
    def match_device(self, pairs, only_ok: bool = True):
        """Run K1..K4 + packing for every chunk of the pair list; results stay on
        the device.  Returns the list of per-chunk dicts."""
        pairs = np.ascontiguousarray(np.asarray(pairs, np.int32)).reshape(-1, 2)
        return [self._match_chunk(pairs[lo:hi], only_ok) for lo, hi in self._chunks(pairs)]

    def _mark(self, name):
        """Stage boundary for ``stage_times``: a CUDA event on the current stream (only when profiling)."""
        if self._marks is not None:
            ev = self.torch.cuda.Event(enable_timing=True)
            ev.record(self.torch.cuda.current_stream(self.device))
            self._marks.append((name, ev))

    def stage_times(self, pairs, only_ok: bool = True) -> dict:
        """One untimed-by-the-caller pass of ``match_device`` with CUDA events between the stages:
        {'knn', 'filter', 'ransac', 'pack'} -> milliseconds summed over the chunks, plus 'total'
        (first to last event) and 'gap' = total - sum (launch gaps, host metadata staging)."""
        torch = self.torch
        self._marks = []
        try:
            self.match_device(pairs, only_ok)
            torch.cuda.synchronize(self.device)
            marks = self._marks
        finally:
            self._marks = None
        out = {}
        for (_, e0), (name, e1) in zip(marks[:-1], marks[1:]):
            if name != "begin":
                out[name] = out.get(name, 0.0) + e0.elapsed_time(e1)
        out["total"] = marks[0][1].elapsed_time(marks[-1][1])
        out["gap"] = out["total"] - sum(v for k, v in out.items() if k != "total")
        return out

    def _match_chunk(self, pairs, only_ok: bool = True):
        """One chunk: dict of device tensors (the kNN / match arenas are reusable buffers)."""
        torch, ctx = self.torch, self.ctx
        P = len(pairs)
        ro = ctx.query_rows(pairs)
        n = int(ro[-1])
        knn = self._buf("knn", (n, 4), torch.int32)
        self._mark("begin")
        ctx.knn2(pairs, ro, knn)
        knn_rev = ro_rev = None
        sparse = self.mutual and not self._full_reverse()
        if self.mutual and not sparse:
            rev = np.ascontiguousarray(pairs[:, ::-1])
            ro_rev = ctx.query_rows(rev)
            knn_rev = self._buf("knn_rev", (int(ro_rev[-1]), 4), torch.int32)
            ctx.knn2(rev, ro_rev, knn_rev)
        self._mark("knn")
        out = (self._buf("mq", (n,), torch.int32), self._buf("mt", (n,), torch.int32),
               self._buf("md2", (n,), torch.int32), self._buf("count", (P,), torch.int32))
        mq, mt, md2, cnt = ctx.filter_matches(pairs, ro, knn, self.ratio, ro_rev, knn_rev, out=out)
        if sparse:
            # the one host round trip of the sparse mutual check: the per-pair survivor counts size the
            # reverse pass (a few hundred descriptors per pair instead of the whole train image)
            if getattr(self, "_cnt_pin", None) is None or self._cnt_pin.numel() < P:
                self._cnt_pin = torch.empty(max(P, 1024), dtype=torch.int32, pin_memory=True)
            self._cnt_pin[:P].copy_(cnt, non_blocking=True)
            torch.cuda.current_stream(self.device).synchronize()
            ctx.mutual_filter(pairs, ro, self._cnt_pin[:P].numpy(), mq, mt, md2, cnt)
        self._mark("filter")
        v = ctx.verify_pairs(pairs, ro, mq, mt, cnt, self.params)
        self._mark("ransac")
        pack_off, dense = ctx.pack_results(ro, cnt, v["status"], only_ok, mq=mq, mt=mt, md2=md2,
                                           mask=v["mask"], sym=v["sym"], pts1=v["pts1"], pts2=v["pts2"])
        self._mark("pack")
        return dict(pairs=pairs, row_off=ro, knn=knn, count=cnt, status=v["status"], metrics=v["metrics"],
                    F=v["F"], pack_off=pack_off, dense=dense)

    def match_pairs_device(self, pairs, only_ok: bool = True) -> dict:
        """Like match_pairs but the results stay on the device, in the layout
        dist.gather_batch_results expects (per-pair tensors + 'kept' + dense arrays): the
        multi-GPU path gathers them on rank 0 over NCCL before anything crosses PCIe."""
        torch = self.torch
        pairs = np.ascontiguousarray(np.asarray(pairs, np.int32)).reshape(-1, 2)
        parts = []
        for lo, hi in self._chunks(pairs):
            d = self._match_chunk(pairs[lo:hi], only_ok)
            total = int(d["pack_off"][-1].item())          # one sync per chunk
            bad = torch.from_numpy((self.counts[pairs[lo:hi, 1]] < 2) | (self.counts[pairs[lo:hi, 0]] < 1))
            status = torch.where(bad.to(self.device), torch.full_like(d["status"], L.PAIR_TOO_FEW_FEATURES),
                                 d["status"])
            dn = d["dense"]
            parts.append(dict(
                pairs=torch.from_numpy(pairs[lo:hi]).to(self.device), status=status, count=d["count"].clone(),
                metrics=d["metrics"].clone(), F=d["F"].clone(),
                kept=(d["pack_off"][1:] - d["pack_off"][:-1]).to(torch.int32),
                dense={k: dn[k][:total].clone() for k in ("mq", "mt", "md2", "mask", "sym", "pts1", "pts2")}))
        if len(parts) == 1:
            return parts[0]
        out = {k: torch.cat([q[k] for q in parts]) for k in ("pairs", "status", "count", "metrics", "F", "kept")}
        out["dense"] = {k: torch.cat([q["dense"][k] for q in parts]) for k in parts[0]["dense"]}
        return out

    # ------------------------------------------------------------- host path
    def _to_host(self, tensors: dict) -> dict:
        """Device tensors -> numpy arrays backed by pinned host memory (torch's caching host
        allocator recycles the blocks): every copy is queued asynchronously, one synchronisation."""
        torch = self.torch
        host = {}
        for k, t in tensors.items():
            h = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
            h.copy_(t, non_blocking=True)
            host[k] = h
        torch.cuda.current_stream(self.device).synchronize()
        return {k: h.numpy() for k, h in host.items()}

    def match_pairs(self, pairs, only_ok: bool = True) -> PairBatchResult:
        """Public call: pair list in, host-side verified matches out (chunked)."""
        torch = self.torch
        pairs = np.ascontiguousarray(np.asarray(pairs, np.int32)).reshape(-1, 2)
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(self.device)
        parts = []

        n_chunks = sum(1 for _ in self._chunks(pairs))
        DENSE = (("mq", "query_idx"), ("mt", "train_idx"), ("md2", "sq_dist"), ("mask", "inlier_mask"),
                 ("sym", "sym_err"), ("pts1", "pts1"), ("pts2", "pts2"))
        arena = {}                  # several chunks: the dense arrays land in ONE pinned array each
        state = dict(base=0, pairs_done=0)

        def reserve(dn, total, n_here):
            """Room for `total` more dense entries (of n_here pairs); capacity from the density so far."""
            need = state["base"] + total
            for k, _ in DENSE:
                cur = arena.get(k)
                if cur is not None and cur.shape[0] >= need:
                    continue
                done_p = state["pairs_done"] + n_here
                cap = max(need, int(need * len(pairs) / max(done_p, 1) * 1.15) + 1024)
                new = torch.empty((cap,) + tuple(dn[k].shape[1:]), dtype=dn[k].dtype, pin_memory=True)
                if cur is not None:
                    new[:state["base"]] = cur[:state["base"]]
                arena[k] = new

        def finish(d, done, lo, hi):
            # device -> pinned host on the copy stream, asynchronously; two synchronisations per chunk
            # (the dense arrays' length is only known once pack_off has arrived)
            with torch.cuda.stream(self._copy_stream):
                self._copy_stream.wait_event(done)
                small = self._to_host(dict(status=d["status"], count=d["count"], metrics=d["metrics"], F=d["F"],
                                           pack_off=d["pack_off"]))
                total = int(small["pack_off"][-1])
                dn = d["dense"]
                if n_chunks == 1:
                    dense = self._to_host({k: dn[k][:total] for k, _ in DENSE})
                else:
                    reserve(dn, total, hi - lo)
                    b = state["base"]
                    for k, _ in DENSE:
                        arena[k][b:b + total].copy_(dn[k][:total], non_blocking=True)
                    self._copy_stream.synchronize()
                    dense = None
            nt_small = self.counts[pairs[lo:hi, 1]] < 2
            nq_zero = self.counts[pairs[lo:hi, 0]] < 1
            # reference gate :268 and the Nt<2 failure of the ratio loop (:151)
            status = np.where(nt_small | nq_zero, L.PAIR_TOO_FEW_FEATURES, small["status"]).astype(np.int32)
            part = dict(pairs=pairs[lo:hi], status=status, count=small["count"],
                        metrics=L.metrics_to_numpy(small["metrics"]), F=small["F"].reshape(-1, 3, 3),
                        pack_off=small["pack_off"])
            if dense is not None:
                part.update({name: dense[k] for k, name in DENSE})
                part["inlier_mask"] = part["inlier_mask"].view(np.bool_)
            parts.append(part)
            state["base"] += total
            state["pairs_done"] += hi - lo

        # chunk c's kernels are queued before chunk c-1's results are fetched: the copies (and this
        # thread's waits) overlap the next chunk's compute.  Everything fetched is a per-chunk
        # allocation except the count arena, which is cloned.
        pending = None
        for lo, hi in self._chunks(pairs):
            d = self._match_chunk(pairs[lo:hi], only_ok)
            d["count"] = d["count"].clone()
            done = torch.cuda.Event()
            done.record(torch.cuda.current_stream(self.device))
            if pending is not None:
                finish(*pending)
            pending = (d, done, lo, hi)
        if pending is not None:
            finish(*pending)
        dense_all = None
        if n_chunks > 1:
            dense_all = {name: arena[k][:state["base"]].numpy() for k, name in DENSE}
            dense_all["inlier_mask"] = dense_all["inlier_mask"].view(np.bool_)
        return _concat(parts, pairs, dense_all)


def _concat(parts, pairs, dense_all=None) -> PairBatchResult:
    """dense_all: the dense arrays of all parts, already contiguous (match_pairs with several chunks)."""
    if not parts:
        z = np.zeros(0, np.int32)
        return PairBatchResult(pairs=pairs, status=z, count=z, metrics=np.zeros(0, L.METRICS_DTYPE),
                               F=np.zeros((0, 3, 3)), pack_off=np.zeros(1, np.int64), query_idx=z,
                               train_idx=z, sq_dist=z, inlier_mask=np.zeros(0, bool),
                               sym_err=np.zeros(0, np.float32), pts1=np.zeros((0, 2), np.float32),
                               pts2=np.zeros((0, 2), np.float32))
    if len(parts) == 1:
        q = parts[0]
        return PairBatchResult(pairs=pairs, **{k: q[k] for k in (
            "status", "count", "metrics", "F", "pack_off", "query_idx", "train_idx", "sq_dist", "inlier_mask",
            "sym_err", "pts1", "pts2")})
    offs, base = [np.zeros(1, np.int64)], 0
    for q in parts:
        offs.append(q["pack_off"][1:] + base)
        base += int(q["pack_off"][-1])
    cat = lambda k: dense_all[k] if dense_all is not None and k in dense_all else \
        np.concatenate([q[k] for q in parts])   # noqa: E731
    return PairBatchResult(
        pairs=pairs, status=cat("status"), count=cat("count"), metrics=cat("metrics"), F=cat("F"),
        pack_off=np.concatenate(offs), query_idx=cat("query_idx"), train_idx=cat("train_idx"),
        sq_dist=cat("sq_dist"), inlier_mask=cat("inlier_mask"), sym_err=cat("sym_err"),
        pts1=cat("pts1"), pts2=cat("pts2"))
