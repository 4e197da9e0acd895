"""Checksums (sum, sum of squares, max) of all 136 fp32-mode projections against the reference's golden values."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from deconv_api_b200 import deepdream as dd, synth, vgg16
g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "vgg16_block5_conv3_reference.npz"))
model = vgg16.VGG16(weights=None, include_top=True)
res = dd.visualize_all_layers(model, synth.image_batch(1), "block5_conv3", "all", precision="fp32")
for name in res:
    st = np.array([[a.astype(np.float64).sum(), (a.astype(np.float64) ** 2).sum(), a.max()] for a in res[name]])
    ref = g["stats|%s" % name]
    rel = np.abs(st - ref) / np.maximum(np.abs(ref), 1e-300)
    print("%-14s max rel err  sum %.2e  sumsq %.2e  max %.2e" % (name, rel[:, 0].max(), rel[:, 1].max(), rel[:, 2].max()))
