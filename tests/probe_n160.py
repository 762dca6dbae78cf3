"""MMA issue-pattern probe for a ring of THREE 160-column accumulators against the wide kernel's two of 256
(B200 box; not a pytest module).  TOP/s incl. the fold K-step (5 K-steps per accumulator)."""
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sfm_b200 import lib as L  # noqa: E402

ctx = L.Context(0)
out = {}
for name, bit, scale in (("N128x2", 0, 1.0), ("N256", 64, 1.0), ("N160_ring3", 128, 160.0 / 256.0)):
    for tag, fl, cons in (("mma_only", 0, 0), ("commit", 1, 0), ("handshake_8warps", 3, 8), ("handshake_4warps", 3, 4)):
        vals = [round(ctx.mma_handshake_probe(5, bit | fl, cons) * scale, 1) for _ in range(3)]
        out[f"{name}_{tag}"] = vals
        print(name, tag, vals, flush=True)
Path("gpurun_out").mkdir(exist_ok=True)
Path("gpurun_out/r02_probe_n160.json").write_text(json.dumps(out, indent=1))
