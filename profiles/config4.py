"""BASELINE configs[3] on one GPU's share: 8 images (64 images / 8 GPUs), for each the forward tape to block5_conv3 and
the projection of all 512 feature maps, bf16 chain, results left on the device (19.7 GB fp32 for the full config would
not be gathered per request)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from deconv_api_b200 import synth, vgg16

model = vgg16.VGG16(weights=None, include_top=False, input_shape=(224, 224, 3))
eng = model.bound_engine(0, upto="block5_conv3")
li = eng.index["block5_conv3"]
feats = np.arange(512, dtype=np.int32)
out = torch.empty(512 * 224 * 224 * 3, dtype=torch.float32, device="cuda")
imgs = [synth.image_batch(1, seed0=s) for s in range(8)]
def run():
    for im in imgs:
        eng.forward(im, li)
        eng.project_dev(li, feats, out.data_ptr(), precision="bf16")
    eng.sync()
run()
t = time.perf_counter(); run(); dt = time.perf_counter() - t
print("8 images x 512 block5_conv3 maps: %.1f ms, %.0f projections/s per GPU (config 4 = 8 GPUs x this, no exchange)" % (dt * 1e3, 8 * 512 / dt))
