"""Writes the preprocess outputs of the 36 bunny images (features from tests/golden/bunny_sift.npz)
with the B200 path into a directory: the input of tests/run_reference_reconstruct.py.
    python tests/make_bunny_outputs.py OUT_DIR        (needs a GPU)"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from sfm_b200 import find_matches as fmx  # noqa: E402


def main():
    out = Path(sys.argv[1])
    out.mkdir(parents=True, exist_ok=True)
    z = np.load(ROOT / "tests" / "golden" / "bunny_sift.npz")
    off, ids = z["offsets"], z["ids"]
    feats = {int(i): (out / "images" / f"{int(i):04d}.ppm", z["kps"][off[n]:off[n + 1]], z["desc"][off[n]:off[n + 1]])
             for n, i in enumerate(ids)}
    m = fmx.ImageMatcher(out, write_visualizations=False)
    m.process_features(feats, int(ids.min()), int(ids.max()))
    m.save_results(out / "matching_results.csv")
    print("pairs passing:", len(m.results))


if __name__ == "__main__":
    main()
