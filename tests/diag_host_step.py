"""Diagnostic (B200 box): host enqueue time against device time of one rank's share of the 200 x 8k workload
at N = 8 (2,487 pairs), per call of PairEngine.match_device, with a cProfile of the host side."""
import cProfile, io, json, pstats, sys, time
from pathlib import Path
import numpy as np, torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sfm_b200 import synth  # noqa: E402
from sfm_b200.engine import PairEngine  # noqa: E402
s = synth.make_image_set(200, 8192, seed=synth.BASE_SEED + 2, tie_frac=0.01)
eng = PairEngine(0)
eng.upload_images(s.desc, s.kps, s.offsets.numpy())
pairs = synth.all_pairs(200).numpy()[:2487]
for _ in range(3):
    eng.match_device(pairs)
torch.cuda.synchronize()
host, dev = [], []
for _ in range(10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e0.record()
    eng.match_device(pairs)
    e1.record()
    host.append((time.perf_counter() - t0) * 1e3)
    torch.cuda.synchronize()
    dev.append(e0.elapsed_time(e1))
pr = cProfile.Profile()
pr.enable()
for _ in range(5):
    eng.match_device(pairs)
pr.disable()
torch.cuda.synchronize()
st = io.StringIO()
pstats.Stats(pr, stream=st).sort_stats("cumulative").print_stats(18)
out = {"pairs": len(pairs), "host_enqueue_ms": [round(x, 2) for x in host], "device_ms": [round(x, 2) for x in dev]}
print(json.dumps(out))
print(st.getvalue()[:5000])
