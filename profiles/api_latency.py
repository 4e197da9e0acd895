"""Latency of the drop-in calls on one B200 (wall clock, host arrays in and out):
  * visualize_all_layers(model, img, 'block5_conv3')  -> 17 layers x top-8 = 136 projections (what reference main.py runs)
  * the same restricted to the requested layer (additive `only_layers` knob)
  * the fused request tail (dcv_request_mosaic, DCV_API_FAST=1): uint8 image in (150 KB), uint8 2x2 mosaic out (602 KB)
  * BASELINE configs[0]: forward to block5_conv3 + ONE projection of the strongest map."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from deconv_api_b200 import deepdream as dd, synth, vgg16

model = vgg16.VGG16(weights=None, include_top=False, input_shape=(224, 224, 3))
img = synth.image_batch(1)
def best(fn, n=5):
    fn(); ts = []
    for _ in range(n):
        t = time.perf_counter(); fn(); ts.append(time.perf_counter() - t)
    return min(ts) * 1e3
for prec in ("fp32", "bf16"):
    full = best(lambda: dd.visualize_all_layers(model, img, "block5_conv3", "all", precision=prec))
    one = best(lambda: dd.visualize_all_layers(model, img, "block5_conv3", "all", precision=prec, only_layers=["block5_conv3"]))
    eng = model.bound_engine(upto="block5_conv3"); li = eng.index["block5_conv3"]
    def cfg1():
        eng.forward(img, li)
        f = int(np.argmax(eng.channel_sums(li)))
        eng.project(li, [f], precision=prec)
    c1 = best(cfg1, 10)
    u8 = synth.smooth_image_u8(5, 224, 224)[..., ::-1].copy()
    l0 = eng.launch_count
    eng.request_mosaic(u8, li, precision=prec)
    launches = eng.launch_count - l0
    tail = best(lambda: eng.request_mosaic(u8, li, precision=prec), 10)
    print("%s: visualize_all_layers full sweep (136 projections) %.1f ms = %.0f proj/s | requested layer only (8) %.1f ms | "
          "fused request tail (preprocess + forward + top-4 + 4 projections + mosaic + de-process, %d launches, H2D 150528 B, "
          "D2H 602112 B + 20 B) %.2f ms | config 1 (forward + 1 projection) %.2f ms = %.0f proj/s"
          % (prec, full, 136e3 / full, one, launches, tail, c1, 1e3 / c1))
