"""Multi-GPU acceptance of the reference-facing entry point: the same synthetic image
directory is preprocessed by one process and by N ranks under torchrun; every output file and the
CSV must be identical.  Needs N GPUs:  python tests/run_ranks_compare.py [N] [n_images]
Prints one JSON line (also the wall times of both runs)."""
import json
import shutil
import subprocess
import sys
import tempfile
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent


def make_dataset(root: Path, n_images: int, seed=7):
    """Textured scene seen under small affine motions: enough SIFT matches for the gates to pass."""
    import cv2
    rng = np.random.default_rng(seed)
    h, w = 480, 640
    base = np.zeros((h * 2, w * 2), np.float32)
    for _ in range(1500):
        c = (int(rng.integers(0, w * 2)), int(rng.integers(0, h * 2)))
        cv2.circle(base, c, int(rng.integers(3, 14)), float(rng.uniform(40, 255)), -1)
    base = cv2.GaussianBlur(base, (0, 0), 1.2)
    (root / "images").mkdir(parents=True)
    for i in range(n_images):
        ang = 3.0 * i
        M = cv2.getRotationMatrix2D((w, h), ang, 1.0 + 0.02 * i)
        M[:, 2] += (-w / 2 + 6 * i, -h / 2 - 4 * i)
        img = cv2.warpAffine(base, M, (w, h)).clip(0, 255).astype(np.uint8)
        cv2.imwrite(str(root / "images" / f"{i:04d}.ppm"), cv2.cvtColor(img, cv2.COLOR_GRAY2BGR))


def run(data_dir: Path, n_images: int, nproc: int):
    cmd = [sys.executable]
    if nproc > 1:
        cmd += ["-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr", "127.0.0.1",
                "--master-port", "29533"]
    cmd += ["-m", "sfm_b200.preprocess", "--data_dir", str(data_dir), "--start_idx", "0", "--end_idx", str(n_images - 1)]
    t0 = time.perf_counter()
    p = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    if p.returncode != 0:
        raise SystemExit(f"{cmd} failed:\n{p.stdout[-2000:]}\n{p.stderr[-4000:]}")
    return time.perf_counter() - t0


def same_tree(a: Path, b: Path):
    fa = sorted(p.relative_to(a) for p in a.rglob("*") if p.is_file() and "images" not in p.parts)
    fb = sorted(p.relative_to(b) for p in b.rglob("*") if p.is_file() and "images" not in p.parts)
    assert fa == fb, (set(fa) ^ set(fb))
    for rel in fa:
        if rel.suffix == ".npz":
            za, zb = np.load(a / rel), np.load(b / rel)
            assert za.files == zb.files, rel
            for k in za.files:
                assert za[k].dtype == zb[k].dtype and np.array_equal(za[k], zb[k], equal_nan=True), (rel, k)
        elif rel.suffix == ".npy":
            xa, xb = np.load(a / rel), np.load(b / rel)
            assert xa.dtype == xb.dtype and np.array_equal(xa, xb), rel
        else:
            assert (a / rel).read_bytes() == (b / rel).read_bytes(), rel
    return len(fa)


def main():
    nproc = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    n_images = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    tmp = Path(tempfile.mkdtemp(prefix="sfm_ranks_"))
    try:
        one, many = tmp / "one", tmp / "many"
        make_dataset(one, n_images)
        shutil.copytree(one, many)
        t1 = run(one, n_images, 1)
        tn = run(many, n_images, nproc)
        n_files = same_tree(one, many)
        rows = (one / "matching_results.csv").read_text().count("\n") - 1
        print(json.dumps({"check": "preprocess outputs identical", "ranks": nproc, "images": n_images,
                          "pairs": n_images * (n_images - 1) // 2, "passing_pairs": rows, "files_compared": n_files,
                          "wall_s_1_rank": round(t1, 2), f"wall_s_{nproc}_ranks": round(tn, 2)}))
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    main()
