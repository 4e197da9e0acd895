"""The reference-faithful CPU baseline of BASELINE.md §3(1): reference app/deepdream.py run UNMODIFIED (its 4-deep Python
max-pool loop and per-channel np.kron unpool included) on the Keras shim, whole visualize_all_layers call.

Runs only where /root/reference exists (the build container; the GPU box has no copy and bench.py may not read it), so
the result is committed as profiles/r2_reference_faithful.json and bench.py reports that record beside its live
`cpu_baseline` (the much stronger vectorised port)."""
import contextlib, io, json, os, platform, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from oracle import keras_shim, reference_loader
from deconv_api_b200 import synth

assert reference_loader.available(), "needs /root/reference"
ref = reference_loader.load_deepdream()
cores = os.cpu_count()
torch.set_num_threads(cores)
model = keras_shim.VGG16(include_top=True)
img = synth.image_batch(1)
rows = {}
for layer in ("block5_conv3", "block4_conv3"):
    best, nproj = None, 0
    for rep in range(2):
        t0 = time.perf_counter()
        with contextlib.redirect_stdout(io.StringIO()):
            out = ref.visualize_all_layers(model, img.copy(), layer, "all")
        dt = time.perf_counter() - t0
        nproj = sum(len(v) for v in out.values())
        best = dt if best is None else min(best, dt)
    rows[layer] = {"seconds_per_request": best, "projections": nproj, "projections_per_s": nproj / best}
    print(layer, rows[layer])
cpu = ""
try:
    cpu = [l.split(":")[1].strip() for l in open("/proc/cpuinfo") if l.startswith("model name")][0]
except Exception:
    cpu = platform.processor()
rec = {"kind": "reference", "what": "reference app/deepdream.py unmodified on oracle/keras_shim (conv arithmetic = torch-CPU fp32; TensorFlow 1.13 "
       "unavailable on this box): visualize_all_layers(model, image, 'block5_conv3', 'all') = 17 layers x top-8 = 136 projections",
       "value": rows["block5_conv3"]["projections_per_s"], "unit": "projections/s", "cores": cores, "cpu": cpu,
       "where": "build container (no GPU); measured by profiles/reference_faithful.py, NOT by the bench run that prints this",
       "requests": rows}
json.dump(rec, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "r2_reference_faithful.json"), "w"), indent=1)
print(json.dumps(rec))
