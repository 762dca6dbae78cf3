"""tcgen05.commit / mbarrier hand-over latencies (B200 box; not a pytest module)."""
import json, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sfm_b200 import lib as L  # noqa: E402
ctx = L.Context(0)
out = {}
for n in (128, 256):
    for k in (0, 1, 2, 5, 10):
        out[f"N{n}_mma{k}"] = ctx.handover_probe(k, n)
for tr in (0, 5, 10):
    out[f"N128_mma5_trail{tr}"] = ctx.handover_probe(5, 128, trail=tr)
print(json.dumps(out, indent=1))
Path("gpurun_out").mkdir(exist_ok=True)
Path("gpurun_out/r02_probe_handover.json").write_text(json.dumps(out, indent=1))
