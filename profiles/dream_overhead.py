"""Fixed cost of one dcv_dream call at 512x512 / 3 octaves: wall time against the number of iterations per octave."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deconv_api_b200 import dream as dm, synth, vgg16
S = int(sys.argv[1]) if len(sys.argv) > 1 else 512
model = vgg16.VGG16(weights=None, include_top=False, input_shape=(None, None, 3))
eng = model.bound_engine(0, upto="block5_conv3")
idx = [eng.index[n] for n in dm.DEFAULT_LAYERS]
coeffs = list(dm.DEFAULT_LAYERS.values())
img = synth.caffe_preprocess(synth.smooth_image_u8(0, S, S))
shapes = dm.octave_shapes(S, S, 3, 1.4)
for iters in (20, 20, 10, 2, 1):
    jt = dm.jitter_offsets(iters * len(shapes), 32, 0)
    eng.dream(img, idx, coeffs, shapes, iters, 1.0, jt, precision="bf16")
    ts = []
    for _ in range(5):
        t0 = time.perf_counter()
        eng.dream(img, idx, coeffs, shapes, iters, 1.0, jt, precision="bf16")
        ts.append(time.perf_counter() - t0)
    print("%2d iterations per octave: %.3f ms per dream (min of 5)" % (iters, min(ts) * 1e3))
