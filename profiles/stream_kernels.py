"""HBM-bound kernels at sizes that are not launch-latency-bound (>= 50 MB per launch), timed with
the CUDA events the library records around each kernel (host<->device copies of the standalone entry points are
outside the events).  Prints achieved GB/s of ALGORITHMIC bytes and the fraction of the measured HBM peak."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from deconv_api_b200 import engine as E

peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
    os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else 6447.5
eng = E.Engine(0)
rng = np.random.default_rng(0)

def run(name, fn, reps=3):
    fn()
    eng.profile(True)
    for _ in range(reps):
        fn()
    rows = eng.profile_launches()
    eng.profile(False)
    ms = sum(r[1] for r in rows) / reps
    by = sum(r[3] for r in rows) / reps
    print("%-34s %8.3f ms  %8.1f MB algorithmic  %7.1f GB/s  %5.1f %% of %.0f" % (name, ms, by / 1e6, by / ms / 1e6, 100 * by / ms / 1e6 / peak, peak))

x = np.maximum(rng.normal(0, 1, (16, 224, 224, 64)), 0).astype(np.float32)           # 205 MB
pooled, sw = eng.maxpool_switch(x)
run("maxpool_switch 16x224x224x64", lambda: eng.maxpool_switch(x))
run("unpool_switch  -> 16x224x224x64", lambda: eng.unpool_switch(pooled, sw))
y = rng.normal(0, 1, (448 * 448 * 3 * 64,)).astype(np.float32)                       # 154 MB
run("deprocess 38.5 M elements", lambda: eng.deprocess(y, mutate=False))
img = rng.normal(0, 1, (8, 1462, 1462, 3)).astype(np.float32)
run("resize_bilinear 8x1462^2 -> 2048^2", lambda: eng.resize_bilinear(img, (2048, 2048)))
big = rng.normal(0, 1, (8, 2048, 2048, 3)).astype(np.float32)
run("roll 8x2048x2048x3", lambda: eng.roll(big, 17, -29))
