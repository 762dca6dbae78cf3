"""TMEM load latency / throughput by shape and warp count (B200 box; not a pytest module):
cycles per tcgen05.ld of one warp and bytes per cycle per SM."""
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sfm_b200 import lib as L  # noqa: E402

ctx = L.Context(0)
names = {0: "x32_wait_each", 1: "x32_wait_every_4", 2: "x64_wait_each", 4: "x16_wait_each", 5: "x16_wait_every_2",
         6: "x16_wait_every_8"}
out = {}
iters = 4096
for v, name in names.items():
    for warps in (1, 2, 4, 8, 16):
        bpc, cyc = ctx.tmem_ld_probe(warps, v, iters)
        out[f"{name}_w{warps}"] = {"cycles_per_load": round(cyc / iters, 1), "bytes_per_cycle_per_sm": round(bpc, 1)}
        print(name, warps, out[f"{name}_w{warps}"], flush=True)
Path("gpurun_out").mkdir(exist_ok=True)
Path("gpurun_out/r02_probe_tmem_ld.json").write_text(json.dumps(out, indent=1))
