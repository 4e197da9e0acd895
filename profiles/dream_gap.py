"""How much of a graph-replayed DeepDream iteration is gaps between kernels: wall time per iteration of a
single-octave run (CUDA graph replay) against the sum of the kernels' own durations (ncu, warm caches).
    python profiles/dream_gap.py [size] [iters]          # prints wall us / iteration
    ncu --metrics gpu__time_duration.sum --cache-control none --clock-control none --csv ... python profiles/dream_gap.py 512 3"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deconv_api_b200 import dream as dm, synth, vgg16
S = int(sys.argv[1]) if len(sys.argv) > 1 else 512
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 200
model = vgg16.VGG16(weights=None, include_top=False, input_shape=(None, None, 3))
eng = model.bound_engine(0, upto="block5_conv3")
idx = [eng.index[n] for n in dm.DEFAULT_LAYERS]
coeffs = list(dm.DEFAULT_LAYERS.values())
jt = dm.jitter_offsets(iters, 32, 0)
img = synth.caffe_preprocess(synth.smooth_image_u8(0, S, S))
eng.dream(img, idx, coeffs, [(S, S)], iters, 1.0, jt, precision="bf16")
t0 = time.perf_counter()
eng.dream(img, idx, coeffs, [(S, S)], iters, 1.0, jt, precision="bf16")
dt = time.perf_counter() - t0
print("%dx%d: %.1f us wall per iteration (%d iterations, one octave, H2D/D2H of the image included once)" % (S, S, dt / iters * 1e6, iters))
