"""``main.py preprocess`` on the B200 path (main.py:189-213, argument names of main.py:11-60).

    python -m sfm_b200.preprocess --data_dir D --start_idx 0 --end_idx 35 [--mask]
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 -m sfm_b200.preprocess ...

One process per GPU under torchrun; every rank writes the files of its own pairs, rank 0 writes
``matching_results.csv``.
"""
from __future__ import annotations

import argparse
import logging
import sys

from .find_matches import run_preprocessing


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    ap.add_argument("--data_dir", required=True)
    ap.add_argument("--start_idx", type=int, default=0)
    ap.add_argument("--end_idx", type=int, default=35)
    ap.add_argument("--mask", action="store_true")
    a = ap.parse_args(argv)
    logging.basicConfig(level=logging.INFO, format="%(asctime)s - %(levelname)s - %(message)s")
    ok = run_preprocessing(a.data_dir, a.start_idx, a.end_idx, mask=a.mask or None)
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
