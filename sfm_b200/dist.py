"""Multi-GPU plumbing: one process per GPU over torch.distributed.

The hot path shards by image PAIR: pairs are independent, so the
only exchanges are (1) replicating the descriptor set on every GPU once (NCCL
all-gather over NVLink) and (2) collecting results: ``ImageMatcher`` lets every rank
write the files of its own pairs (file names are per pair, so the directory equals
a 1-GPU run's) and gathers only the CSV rows on rank 0 (``gather_rows``);
``gather_batch_results`` brings whole match lists to rank 0 for callers that want
them in one process.  No collective runs during compute.  Works with the gloo
backend on CPU tensors for tests.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist


def rank_world(group=None):
    """(rank, world) of this process; (0, 1) outside torch.distributed."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def all_gather_objects(obj, group=None):
    """Small picklable host objects from every rank, in rank order, on every rank."""
    out = [None] * dist.get_world_size(group)
    dist.all_gather_object(out, obj, group=group)
    return out


def gather_rows(rows: list, group=None):
    """Per-pair result rows (CSV dicts) of every rank concatenated in rank order on rank 0 --
    pair order, because ranks own contiguous ranges of the pair list.  None elsewhere."""
    parts = all_gather_objects(rows, group)
    if dist.get_rank(group) != 0:
        return None
    return [r for part in parts for r in part]


def image_blocks(n_images: int, world: int):
    """Contiguous image ranges per rank (who extracts / uploads which images)."""
    bounds = [(n_images * r) // world for r in range(world + 1)]
    return [(bounds[r], bounds[r + 1]) for r in range(world)]


def partition_pairs(pairs: np.ndarray, counts: np.ndarray, world: int):
    """Split the (i-major) pair list into `world` contiguous ranges of near-equal
    cost, cost(i, j) = n_i * n_j (tensor work of the kNN).  Contiguity keeps one
    query image's tiles hot across consecutive train images on a rank.
    Returns [(lo, hi)] per rank."""
    pairs = np.asarray(pairs).reshape(-1, 2)
    if len(pairs) == 0:
        return [(0, 0)] * world
    cost = counts[pairs[:, 0]].astype(np.float64) * counts[pairs[:, 1]].astype(np.float64) + 1.0
    cum = np.cumsum(cost)
    cuts = [0]
    for r in range(1, world):
        cuts.append(int(np.searchsorted(cum, cum[-1] * r / world, side="left")))
    cuts.append(len(pairs))
    cuts = np.maximum.accumulate(cuts)
    return [(int(cuts[r]), int(cuts[r + 1])) for r in range(world)]


def _gather_sizes(values, device, group=None) -> np.ndarray:
    """int64 vector `values` of every rank as a [world, len(values)] host array: one collective, one host read."""
    world = dist.get_world_size(group)
    mine = torch.tensor([int(v) for v in values], dtype=torch.int64, device=device)
    out = torch.empty((world * mine.numel(),), dtype=torch.int64, device=device)
    if dist.get_backend(group) == "nccl":
        dist.all_gather_into_tensor(out, mine, group=group)
    else:
        parts = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(parts, mine, group=group)
        out = torch.cat(parts)
    return out.cpu().numpy().reshape(world, -1)


def _all_gather_rows(t: torch.Tensor, sizes, group=None):
    """all-gather of tensors whose first dimension is sizes[rank] (known on every rank): the concatenation in
    rank order.  Equal sizes -- the usual case, equal image blocks -- take one ncclAllGather straight into the
    result; ragged sizes are padded to the largest block and the pieces concatenated."""
    world = dist.get_world_size(group)
    sizes = [int(x) for x in sizes]
    mx = max(max(sizes), 1)
    tail = tuple(t.shape[1:])
    nccl = dist.get_backend(group) == "nccl"
    if nccl and min(sizes) == mx:
        out = torch.empty((world * mx,) + tail, dtype=t.dtype, device=t.device)
        dist.all_gather_into_tensor(out, t.contiguous(), group=group)
        return out
    pad = t
    if t.shape[0] != mx:
        pad = torch.zeros((mx,) + tail, dtype=t.dtype, device=t.device)
        pad[: t.shape[0]] = t
    if nccl:
        out = torch.empty((world * mx,) + tail, dtype=t.dtype, device=t.device)
        dist.all_gather_into_tensor(out, pad.contiguous(), group=group)     # one ncclAllGather
        parts = list(out.split(mx))
    else:
        parts = [torch.empty_like(pad) for _ in range(world)]
        dist.all_gather(parts, pad.contiguous(), group=group)
    return torch.cat([p[:n] for p, n in zip(parts, sizes)])


def _all_gather_var(t: torch.Tensor, group=None):
    """all-gather of tensors whose first dimension differs per rank.
    Returns the list of per-rank tensors (on t's device)."""
    sizes = _gather_sizes([t.shape[0]], t.device, group)[:, 0]
    cat = _all_gather_rows(t, sizes, group)
    return list(cat.split([int(n) for n in sizes]))


def replicate_images(local_desc: torch.Tensor, local_kps: torch.Tensor, local_counts, group=None):
    """Every rank contributes the descriptors/keypoints of its contiguous image
    block; returns the full (desc u8 [R,128], kps f32 [R,2], offsets int64 [N+1])
    on every rank, in image order.  Two host reads in all (block sizes, per-image counts); the
    descriptor and keypoint exchanges are enqueued without one."""
    dev = local_desc.device
    local_counts = np.asarray(local_counts, np.int64)
    sizes = _gather_sizes([len(local_counts), local_desc.shape[0]], dev, group)       # [world, 2]
    cnt = torch.as_tensor(local_counts, device=dev)
    counts = _all_gather_rows(cnt, sizes[:, 0], group).cpu().numpy()
    desc = _all_gather_rows(local_desc, sizes[:, 1], group)
    kps = _all_gather_rows(local_kps, sizes[:, 1], group)
    off = np.zeros(len(counts) + 1, np.int64)
    off[1:] = np.cumsum(counts)
    return desc, kps, off


def gather_to_root(t: torch.Tensor, group=None):
    """Variable-length gather of `t` (first dim) on rank 0.  Returns the
    concatenation on rank 0 and None elsewhere.  Padded all-gather: result
    volumes are O(matches), tiny next to NVLink bandwidth."""
    parts = _all_gather_var(t, group)
    if dist.get_rank(group) != 0:
        return None
    return torch.cat(parts)


PER_PAIR_KEYS = ("pairs", "status", "count", "metrics", "F", "kept")


def _pack_bytes(tensors):
    """Concatenate tensors as bytes, every piece padded to 16 bytes (so each can be viewed back)."""
    pieces = []
    for t in tensors:
        b = t.contiguous().view(torch.uint8).reshape(-1)
        pad = (-b.numel()) % 16
        pieces.append(b)
        if pad:
            pieces.append(torch.zeros(pad, dtype=torch.uint8, device=b.device))
    return torch.cat(pieces) if pieces else torch.zeros(0, dtype=torch.uint8)


def _piece_bytes(n_rows, like):
    rb = like.element_size() * int(np.prod(like.shape[1:], dtype=np.int64))
    nb = int(n_rows) * rb
    return nb, nb + (-nb) % 16


def gather_batch_results(local: dict, group=None):
    """local: dict of tensors for this rank's pair range -- per-pair tensors
    (first dim = local pair count) under keys 'pairs','status','count','metrics',
    'F','kept' (number of dense entries per pair) and dense tensors under
    'dense' (dict).  Rank 0 gets the same structure for ALL pairs, in pair order.

    One small all-gather of (pairs, dense entries) per rank sizes everything; then every rank
    sends ONE packed byte buffer to rank 0 and rank 0 posts one receive per rank, all in one
    batch (NCCL: a single group of send/recv) -- no padding to the largest rank, one host
    synchronisation (the count vector)."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    dense_keys = sorted(local["dense"])
    tensors = [local[k] for k in PER_PAIR_KEYS] + [local["dense"][k] for k in dense_keys]
    dev = local["status"].device
    n_dense = int(local["dense"][dense_keys[0]].shape[0]) if dense_keys else 0
    cnt = torch.tensor([int(local["status"].shape[0]), n_dense], dtype=torch.int64, device=dev)
    all_cnt = [torch.zeros_like(cnt) for _ in range(world)]
    dist.all_gather(all_cnt, cnt, group=group)
    all_cnt = torch.stack(all_cnt).cpu().numpy()

    def sizes(r):
        out = []
        for i, t in enumerate(tensors):
            out.append(_piece_bytes(all_cnt[r, 0] if i < len(PER_PAIR_KEYS) else all_cnt[r, 1], t))
        return out

    ops, rbufs = [], {}
    if rank == 0:
        for r in range(1, world):
            total = sum(padded for _, padded in sizes(r))
            rbufs[r] = torch.empty(total, dtype=torch.uint8, device=dev)
            if total:
                ops.append(dist.P2POp(dist.irecv, rbufs[r], r, group))
    else:
        buf = _pack_bytes(tensors)
        if buf.numel():
            ops.append(dist.P2POp(dist.isend, buf, 0, group))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    if rank != 0:
        return None
    parts = [list(tensors)]
    for r in range(1, world):
        o, got = 0, []
        for i, (t, (nb, padded)) in enumerate(zip(tensors, sizes(r))):
            n_rows = all_cnt[r, 0] if i < len(PER_PAIR_KEYS) else all_cnt[r, 1]
            got.append(rbufs[r][o:o + nb].view(t.dtype).reshape((int(n_rows),) + tuple(t.shape[1:])))
            o += padded
        parts.append(got)
    out = {}
    for i, k in enumerate(PER_PAIR_KEYS):
        out[k] = torch.cat([q[i] for q in parts])
    out["dense"] = {k: torch.cat([q[len(PER_PAIR_KEYS) + i] for q in parts]) for i, k in enumerate(dense_keys)}
    kept = out["kept"].to(torch.int64)
    off = torch.zeros(kept.numel() + 1, dtype=torch.int64, device=kept.device)
    off[1:] = torch.cumsum(kept, 0)
    out["pack_off"] = off
    return out
