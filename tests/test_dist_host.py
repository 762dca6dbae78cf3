"""Host-side logic of the multi-GPU path on CPU: pair partitioning, descriptor replication and
result gathering with world_size 2 over gloo, plus the C-ABI surface."""
import ctypes
import os
import re
import socket
from pathlib import Path

import numpy as np
import pytest

torch = pytest.importorskip("torch")
import torch.distributed as dist  # noqa: E402
import torch.multiprocessing as mp  # noqa: E402

from sfm_b200 import dist as D  # noqa: E402

ROOT = Path(__file__).resolve().parent.parent


def test_partition_pairs_balanced_and_complete():
    rng = np.random.default_rng(0)
    counts = rng.integers(52, 500, 36)
    pairs = np.array([(i, j) for i in range(36) for j in range(i + 1, 36)], np.int32)
    cost = counts[pairs[:, 0]] * counts[pairs[:, 1]]
    for world in (1, 2, 4, 8):
        parts = D.partition_pairs(pairs, counts, world)
        assert parts[0][0] == 0 and parts[-1][1] == len(pairs)
        assert all(parts[r][1] == parts[r + 1][0] for r in range(world - 1))     # contiguous, no overlap
        loads = np.array([cost[lo:hi].sum() for lo, hi in parts], np.float64)
        assert loads.max() <= loads.mean() * 1.25 + cost.max()
    assert D.partition_pairs(np.zeros((0, 2), np.int32), counts, 4) == [(0, 0)] * 4
    assert D.image_blocks(10, 4) == [(0, 2), (2, 5), (5, 7), (7, 10)]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(7)
        counts = np.array([5, 0, 9, 3, 7], np.int64)                        # ragged, one empty image
        off = np.zeros(len(counts) + 1, np.int64)
        off[1:] = np.cumsum(counts)
        desc = torch.from_numpy(rng.integers(0, 256, (int(off[-1]), 128)).astype(np.uint8))
        kps = torch.from_numpy(rng.random((int(off[-1]), 2)).astype(np.float32))
        b0, b1 = D.image_blocks(len(counts), world)[rank]
        d, k, o = D.replicate_images(desc[off[b0]:off[b1]], kps[off[b0]:off[b1]], counts[b0:b1])
        ok = torch.equal(d, desc) and torch.equal(k, kps) and np.array_equal(o, off)
        # per-pair results of this rank's share, gathered on rank 0 in pair order
        pairs = np.array([(i, j) for i in range(5) for j in range(i + 1, 5)], np.int32)
        lo, hi = D.partition_pairs(pairs, np.maximum(counts, 1), world)[rank]
        kept = torch.tensor([(int(a) + int(b)) % 4 for a, b in pairs[lo:hi]], dtype=torch.int32)
        dense_q = torch.cat([torch.full((int(n),), 100 * (lo + p), dtype=torch.int32) + torch.arange(int(n), dtype=torch.int32)
                             for p, n in enumerate(kept)] + [torch.zeros(0, dtype=torch.int32)])
        local = dict(pairs=torch.from_numpy(pairs[lo:hi]), status=torch.arange(lo, hi, dtype=torch.int32),
                     count=kept.clone(), metrics=torch.zeros((hi - lo, 32), dtype=torch.uint8),
                     F=torch.arange(lo, hi, dtype=torch.float64)[:, None].repeat(1, 9), kept=kept,
                     dense=dict(mq=dense_q))
        g = D.gather_batch_results(local)
        if rank == 0:
            all_kept = np.array([(int(a) + int(b)) % 4 for a, b in pairs])
            ok = ok and torch.equal(g["pairs"], torch.from_numpy(pairs))
            ok = ok and torch.equal(g["status"], torch.arange(len(pairs), dtype=torch.int32))
            ok = ok and np.array_equal(g["pack_off"].numpy(), np.concatenate([[0], np.cumsum(all_kept)]))
            want = np.concatenate([100 * p + np.arange(n) for p, n in enumerate(all_kept)]).astype(np.int32)
            ok = ok and np.array_equal(g["dense"]["mq"].numpy(), want)
            ok = ok and torch.equal(g["F"][:, 0], torch.arange(len(pairs), dtype=torch.float64))
        else:
            ok = ok and g is None
        # CSV rows of ImageMatcher's rank path: pair order on rank 0, a rank may have none
        ok = ok and D.rank_world() == (rank, world)
        rows = [{"pair_name": f"pair_{a}_{b}", "num_matches": int(a) * 10 + int(b)} for a, b in pairs[lo:hi]]
        if rank == 1:
            rows = rows[:0] if hi - lo < 2 else rows
        got = D.gather_rows(rows)
        lists = D.all_gather_objects(sorted(range(b0, b1)))
        ok = ok and [i for part in lists for i in part] == list(range(len(counts)))
        if rank == 0:
            ok = ok and [r["pair_name"] for r in got] == [f"pair_{a}_{b}" for a, b in pairs]
        else:
            ok = ok and got is None
        ret[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_replicate_and_gather_world2_gloo():
    world = 2
    port = _free_port()
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
        assert dict(ret) == {0: True, 1: True}


def test_rank_world_outside_distributed():
    assert D.rank_world() == (0, 1)


def test_c_abi_exports_every_declared_symbol():
    """libsfm_match.so loads and exports exactly the functions include/sfm_match.h declares
    (no compute call: there is no GPU here)."""
    from sfm_b200 import build, lib as L
    path = build.build()
    so = ctypes.CDLL(str(path))
    header = (ROOT / "include" / "sfm_match.h").read_text()
    declared = set(re.findall(r"SFM_API\s+[\w\s\*]+?\b(sfm_\w+)\s*\(", header))
    assert declared and declared == set(L.EXPORTS)
    for name in declared:
        assert hasattr(so, name), name
    assert so.sfm_abi_version() == L.ABI_VERSION
    so.sfm_last_error.restype = ctypes.c_char_p
    h = ctypes.c_void_p()
    if not torch.cuda.is_available():
        assert so.sfm_create(0, ctypes.byref(h)) != 0          # fails loudly without a device
        assert b"CUDA" in so.sfm_last_error(None) or b"device" in so.sfm_last_error(None)


@pytest.mark.gpu
@pytest.mark.timeout(900)
def test_preprocess_two_ranks_equal_one_rank():
    """`python -m sfm_b200.preprocess` under torchrun with 2 ranks writes the same files and CSV as one
    process (tests/run_ranks_compare.py); needs two GPUs."""
    import json
    import subprocess
    import sys
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    p = subprocess.run([sys.executable, str(ROOT / "tests" / "run_ranks_compare.py"), "2", "8"], capture_output=True,
                       text=True, timeout=800)
    assert p.returncode == 0, p.stderr[-3000:]
    out = json.loads(p.stdout.strip().splitlines()[-1])
    assert out["check"] == "preprocess outputs identical" and out["files_compared"] > 20
