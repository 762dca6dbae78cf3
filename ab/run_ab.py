import runpy, sys
from pathlib import Path
import sfm_b200.lib as L
if sys.argv[1] == "old":
    L.LIB_PATH = Path(__file__).resolve().parent / "libsfm_old.so"
sys.argv = ["diag_wide.py", "36"]
runpy.run_path(str(Path(__file__).resolve().parent.parent / "tests" / "diag_wide.py"), run_name="__main__")
