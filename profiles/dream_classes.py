"""Per-kernel-class device time of one DeepDream run (CUDA events inside the library).
    python profiles/dream_classes.py [bf16|fp32] [size]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from deconv_api_b200 import dream as dm, synth, vgg16

prec = sys.argv[1] if len(sys.argv) > 1 else "bf16"
S = int(sys.argv[2]) if len(sys.argv) > 2 else 512
model = vgg16.VGG16(weights=None, include_top=False, input_shape=(None, None, 3))
eng = model.bound_engine(0, upto="block5_conv3")
idx = [eng.index[n] for n in dm.DEFAULT_LAYERS]
coeffs = list(dm.DEFAULT_LAYERS.values())
shapes = dm.octave_shapes(S, S)
jt = dm.jitter_offsets(60, 32, 0)
img = synth.caffe_preprocess(synth.smooth_image_u8(0, S, S))
import time
eng.dream(img, idx, coeffs, shapes, 20, 1.0, jt, precision=prec)
t0 = time.perf_counter(); eng.dream(img, idx, coeffs, shapes, 20, 1.0, jt, precision=prec); wall = (time.perf_counter() - t0) * 1e3
eng.profile(True)
eng.dream(img, idx, coeffs, shapes[-1:], 1, 1.0, jt[:1], precision=prec)       # ONE iteration at full size
rows = eng.profile_launches()
eng.profile(False)
tot = sum(r[1] for r in rows)
print("wall per dream %.2f ms (%.1f iters/s); one %dx%d iteration: %d profiled launches, %.3f ms of kernels" % (wall, 60e3 / wall, S, S, len(rows), tot))
for i, (c, ms, fl, by) in enumerate(rows):
    print("%-3d %-10s %8.4f ms %9.1f TFLOP/s %9.1f GB/s" % (i, c, ms, fl / ms / 1e9 if ms else 0, by / ms / 1e6 if ms else 0))
