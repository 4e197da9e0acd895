"""Per-launch device times of one fused API request (dcv_request_mosaic, block5_conv3) and its wall time."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deconv_api_b200 import synth, vgg16
prec = sys.argv[1] if len(sys.argv) > 1 else "bf16"
model = vgg16.VGG16(weights=None, include_top=False, input_shape=(224, 224, 3))
eng = model.bound_engine(upto="block5_conv3"); li = eng.index["block5_conv3"]
u8 = synth.smooth_image_u8(5, 224, 224)[..., ::-1].copy()
for _ in range(3): eng.request_mosaic(u8, li, precision=prec)
ts = []
for _ in range(10):
    t = time.perf_counter(); eng.request_mosaic(u8, li, precision=prec); ts.append(time.perf_counter() - t)
print("wall %.3f ms (min of 10)" % (min(ts) * 1e3))
eng.profile(True)
eng.request_mosaic(u8, li, precision=prec)
rows = eng.profile_launches()
eng.profile(False)
tot = sum(r[1] for r in rows)
for i, (c, ms, fl, by) in enumerate(rows):
    print("%-3d %-10s %8.4f ms" % (i, c, ms))
print("sum of launches %.3f ms over %d" % (tot, len(rows)))
