"""Diagnostic: max-pool switches of the library's forward against the CPU oracle on a 96x160 image with bias-carrying
random weights; for every mismatch print how close the two candidates of the pooling window are."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import torch.nn.functional as F
from oracle import keras_shim, restatement as rs
from deconv_api_b200 import synth, vgg16
model = vgg16.VGG16(weights=None, include_top=False, input_shape=(96, 160, 3), seed=5, bias=True)
ref = keras_shim.VGG16(include_top=False, input_shape=(96, 160, 3), seed=5, bias=True)
img = synth.caffe_preprocess(synth.smooth_image_u8(3, 96, 160))
eng = model.bound_engine(upto="block3_conv3")
eng.forward(img, eng.index["block3_conv3"])
tape = rs.FastTape(ref, "block3_conv3", img)
pools = [op for op in tape.ops if op[0] == "pool"]
for name, op in zip(("block1_pool", "block2_pool"), pools):
    idx = op[1][0].numpy()                       # (C, h, w) flat indices into the (H, W) input plane
    H, W = op[2]
    dy = (idx // W) % 2; dx = (idx % W) % 2
    sw_ref = (dy * 2 + dx).transpose(1, 2, 0)    # (h, w, C)
    sw = eng.switch_index(eng.index[name])[0]
    bad = np.argwhere(sw != sw_ref)
    print(name, "mismatches:", len(bad), "of", sw.size)
    conv_name = name.replace("pool", "conv2")
    act = eng.activation(eng.index[conv_name])[0]
    for (y, x, c) in bad[:10]:
        win = act[2 * y:2 * y + 2, 2 * x:2 * x + 2, c].ravel()
        print("   window", win, "ours", sw[y, x, c], "oracle", sw_ref[y, x, c], "rel gap %.2e" % (abs(win[sw[y, x, c]] - win[sw_ref[y, x, c]]) / max(abs(win).max(), 1e-30)))
