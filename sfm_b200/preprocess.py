"""``main.py preprocess`` on the B200 path (main.py:189-213; flags of main.py:52-76).

    python -m sfm_b200.preprocess [--log_level L] [preprocess] --data_dir D [--start_idx 0] [--end_idx 35]
                                  [--min_matches 150] [--visualize] [--mask PATH] [--mutual_nn]
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 -m sfm_b200.preprocess ...

The reference's own command line parses unchanged (the ``preprocess`` sub-command word is optional
here).  As in the reference, ``--min_matches`` is validated (20..1000) and otherwise unused, and
``--mask PATH`` must exist and then only switches the per-image silhouette masks on (main.py:199 passes
it for its truthiness).  ``--visualize`` writes ``visualizations/<pair>_matches.png`` for passing pairs
(the reference writes them always, find_matches.py:293-295).  ``--mutual_nn`` adds the mutual
nearest-neighbour check of README.md:108.  One process per GPU under torchrun; every rank writes the
files of its own pairs, rank 0 writes ``matching_results.csv``.  Exit code 0 / 1 as main.py:303-310.
"""
from __future__ import annotations

import argparse
import logging
import sys
from pathlib import Path

from .find_matches import run_preprocessing


def parse_args(argv=None):
    ap = argparse.ArgumentParser(description="preprocess stage (pairwise matching) on the B200 path")
    ap.add_argument("operation", nargs="?", default="preprocess", choices=["preprocess"])
    ap.add_argument("--log_level", default="INFO", choices=["DEBUG", "INFO", "WARNING", "ERROR"])
    ap.add_argument("--data_dir", required=True)
    ap.add_argument("--start_idx", type=int, default=0)
    ap.add_argument("--end_idx", type=int, default=35)
    ap.add_argument("--min_matches", type=int, default=150)
    ap.add_argument("--visualize", action="store_true")
    ap.add_argument("--mask", type=str, default=None)
    ap.add_argument("--mutual_nn", action="store_true")
    a = ap.parse_args(argv)
    for name, (lo, hi) in (("start_idx", (0, 999)), ("end_idx", (0, 999)), ("min_matches", (20, 1000))):
        v = getattr(a, name)                                         # main.py:159-175
        if not lo <= v <= hi:
            ap.error(f"{name} must be between {lo} and {hi}, got {v}")
    if not Path(a.data_dir).exists():                                # main.py:38-46
        ap.error(f"Directory does not exist: {a.data_dir}")
    if a.mask and not Path(a.mask).exists():                         # main.py:177-181
        ap.error(f"Mask file does not exist: {a.mask}")
    return a


def main(argv=None) -> int:
    a = parse_args(argv)
    logging.basicConfig(level=getattr(logging, a.log_level), format="%(asctime)s - %(levelname)s - %(message)s")
    ok = run_preprocessing(a.data_dir, a.start_idx, a.end_idx, mask=a.mask, mutual=a.mutual_nn,
                           visualize=a.visualize)
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
