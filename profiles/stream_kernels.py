"""HBM-bound kernels at sizes that are not launch-latency-bound (>= 50 MB per launch).

Inputs are resident in HBM (torch tensors), each kernel is launched REPS times back to back through the
device-resident C-ABI entry points, and the library's own CUDA events around every launch are read back
(dcv_profile_launch).  The first launch is dropped.  Prints achieved GB/s of ALGORITHMIC bytes and the fraction of
the measured HBM peak.  (Timing a single launch right after a pageable host copy, as the host-buffer entry points
would, measures clock ramp-up: the same roll kernel reads 0.23 ms that way and 0.12 ms back to back.)"""
import ctypes, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deconv_api_b200 import engine as E

REPS = 11
root = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
peak = json.load(open(os.path.join(root, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(root, "MEASURED_PEAKS.json")) else 6447.5
eng = E.Engine(0)
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(0)
results = {}


def run(name, fn):
    torch.cuda.synchronize()
    fn(); eng.sync()
    eng.profile(True)
    for _ in range(REPS):
        fn()
    eng.sync()
    rows = eng.profile_launches()
    eng.profile(False)
    per = len(rows) // REPS                      # launches per call
    rows = rows[per:]                            # drop the first call
    ms = sum(r[1] for r in rows) / (REPS - 1)
    by = sum(r[3] for r in rows) / (REPS - 1)
    gbs = by / ms / 1e6
    results[name] = {"ms": ms, "algorithmic_MB": by / 1e6, "GBps": gbs, "frac_of_peak": gbs / peak}
    print("%-36s %8.3f ms  %8.1f MB algorithmic  %7.1f GB/s  %5.1f %% of %.0f" % (name, ms, by / 1e6, gbs, 100 * gbs / peak, peak))


x = torch.relu(torch.randn((16, 224, 224, 64), device=dev, generator=g))                       # 205 MB
pooled = torch.empty((16, 112, 112, 64), device=dev)
sw = torch.empty((16, 112, 112, 64), device=dev, dtype=torch.uint8)
back = torch.empty_like(x)
run("maxpool_switch 16x224x224x64", lambda: eng.maxpool_switch_dev(x.data_ptr(), x.shape, pooled.data_ptr(), sw.data_ptr()))
run("unpool_switch -> 16x224x224x64", lambda: eng.unpool_switch_dev(pooled.data_ptr(), sw.data_ptr(), x.shape, back.data_ptr()))

y = torch.randn((448 * 448 * 3 * 64,), device=dev, generator=g)                                # 154 MB
y8 = torch.empty(y.shape, device=dev, dtype=torch.uint8)
run("deprocess 38.5 M elements", lambda: eng._ck(eng.lib.dcv_deprocess_dev(eng.ctx, ctypes.c_void_p(y.data_ptr()), y.numel(),
                                                                            ctypes.c_void_p(y8.data_ptr()))))

img = torch.randn((8, 1462, 1462, 3), device=dev, generator=g)
big = torch.empty((8, 2048, 2048, 3), device=dev)
run("resize_bilinear 8x1462^2 -> 2048^2", lambda: eng.resize_bilinear_dev(img.data_ptr(), img.shape, (2048, 2048), big.data_ptr()))
small = torch.empty((8, 1462, 1462, 3), device=dev)
run("resize_bilinear 8x2048^2 -> 1462^2", lambda: eng.resize_bilinear_dev(big.data_ptr(), big.shape, (1462, 1462), small.data_ptr()))
rolled = torch.empty_like(big)
run("roll 8x2048x2048x3", lambda: eng.roll_dev(big.data_ptr(), big.shape, 17, -29, rolled.data_ptr()))

if len(sys.argv) > 1:
    json.dump({"hbm_peak_GBps": peak, "reps": REPS - 1, "kernels": results}, open(sys.argv[1], "w"), indent=1)
