"""BASELINE configs[4] on one GPU: DeepDream on a 2048x2048 synthetic image, 4 octaves x 20 iterations (bf16)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deconv_api_b200 import dream as dm, synth, vgg16
model = vgg16.VGG16(weights=None, include_top=False, input_shape=(None, None, 3))
img = synth.caffe_preprocess(synth.smooth_image_u8(0, 2048, 2048))
print("octaves", dm.octave_shapes(2048, 2048, 4, 1.4))
for prec in ("bf16",):
    dm.dream(model, img, num_octave=4, iterations=20, precision=prec)
    t = time.perf_counter(); out = dm.dream(model, img, num_octave=4, iterations=20, precision=prec); dt = time.perf_counter() - t
    print("%s: %.1f ms per dream, %.1f iterations/s (host image in, host image out)" % (prec, dt * 1e3, 80 / dt))
