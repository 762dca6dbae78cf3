"""Acceptance harness: the UNMODIFIED reference reconstruct stage
(utils/sfm_reconstruction.py: StructureFromMotion.run_reconstruction / save_reconstruction, as
main.py:214-231 drives them) run on a preprocess output directory.

    python tests/run_reference_reconstruct.py DATA_DIR [num_images] [seed]

Runs only where /root/reference exists (the build container, no GPU needed): DATA_DIR is either the
reference's shipped bunny_data or the directory tests/make_bunny_outputs.py wrote with the B200
path on the GPU box.  Environment shims, none of which touches reference code:
  * matplotlib (absent) is a MagicMock, networkx (absent) is tests/stubs/networkx;
  * the CSV path hard-coded at utils/sfm_reconstruction.py:36 and :643 is redirected to
    DATA_DIR/matching_results.csv by wrapping pandas.read_csv;
  * the working directory is a scratch directory (visualize_graph writes image_graph.png to CWD).
Prints one JSON line with what the reference reconstructed."""
import json
import os
import sys
import tempfile
import time
import types
from pathlib import Path
from unittest import mock

import numpy as np

REF = Path("/root/reference")
HARD_CODED = "/teamspace/studios/this_studio/SFM/bunny_data/matching_results.csv"


def main():
    data_dir = Path(sys.argv[1]).resolve()
    num_images = int(sys.argv[2]) if len(sys.argv) > 2 else 36
    seed = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    if not REF.exists():
        raise SystemExit("needs /root/reference")
    sys.path.insert(0, str(Path(__file__).resolve().parent / "stubs"))
    plt = mock.MagicMock()
    plt.subplots.return_value = (mock.MagicMock(), mock.MagicMock())
    mpl = types.ModuleType("matplotlib")
    mpl.pyplot = plt
    sys.modules.setdefault("matplotlib", mpl)
    sys.modules.setdefault("matplotlib.pyplot", plt)
    import pandas as pd
    real_read_csv = pd.read_csv

    def read_csv(path, *a, **kw):
        if str(path) == HARD_CODED:
            path = data_dir / "matching_results.csv"
        return real_read_csv(path, *a, **kw)
    pd.read_csv = read_csv
    sys.path.insert(0, str(REF))
    scratch = tempfile.mkdtemp(prefix="sfm_recon_")
    os.chdir(scratch)
    np.random.seed(seed)
    import cv2
    cv2.setRNGSeed(seed)
    from utils.sfm_reconstruction import StructureFromMotion
    t0 = time.perf_counter()
    sfm = StructureFromMotion(data_dir)
    sfm.run_reconstruction(num_images)
    out = Path(scratch) / "reconstruction"
    sfm.save_reconstruction(out)
    stats = sfm.compute_reconstruction_stats()
    files = sorted(p.name for p in out.iterdir())
    clean = {k: (float(v) if isinstance(v, (int, float, np.floating, np.integer)) else str(v)) for k, v in stats.items()}
    print(json.dumps({"data_dir": str(data_dir), "csv_rows": int(len(real_read_csv(data_dir / "matching_results.csv"))),
                      "images_registered": len(sfm.poses), "registered_ids": sorted(int(i) for i in sfm.poses),
                      "points3D": len(sfm.points3D), "tracks": len(sfm.point_tracks), "stats": clean,
                      "reconstruction_files": files, "seconds": round(time.perf_counter() - t0, 1)}))


if __name__ == "__main__":
    main()
