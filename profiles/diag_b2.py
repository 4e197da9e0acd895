"""Diagnostic: batch-2 VGG16 projection vs the oracle, per (feature, image), and switch agreement per pool layer."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import keras_shim, restatement as rs
from deconv_api_b200 import synth, vgg16

def rel(a, b):
    return float(np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-30))

seed0 = int(sys.argv[1]) if len(sys.argv) > 1 else 3
layer = sys.argv[2] if len(sys.argv) > 2 else "block3_conv2"
model = vgg16.VGG16(weights=None, include_top=False, input_shape=(224, 224, 3))
ref_model = keras_shim.VGG16(include_top=False, input_shape=(224, 224, 3))
img = synth.image_batch(2, 224, 224, seed0)
eng = model.bound_engine(upto=layer)
li = eng.index[layer]
feats = [5, 17, 100, 255, 3, 42]
for tag, data in (("B=2", img), ("img0", img[:1]), ("img1", img[1:])):
    eng.forward(data, li)
    f32 = eng.project(li, feats, precision="fp32")
    for b in range(data.shape[0]):
        gi = b if tag == "B=2" else (0 if tag == "img0" else 1)
        tape = rs.FastTape(ref_model, layer, img[gi:gi + 1])
        ref = tape.project(feats)
        print(tag, "image", gi, ["%.1e" % rel(f32[k, b], ref[k]) for k in range(len(feats))])
    entries = rs.forward(rs.build_entries(ref_model, layer), data)
    for e in entries:
        if e.kind == "pool":
            idx = eng.switch_index(eng.index[e.name])
            prev = entries[entries.index(e) - 1].up
            _, _, ref_idx = rs.maxpool_with_switch(prev)
            mism = np.argwhere(idx != ref_idx)
            print("  ", e.name, "switch mismatches:", len(mism))
            for (b, y, x, c) in mism[:5]:
                patch = prev[b, 2 * y:2 * y + 2, 2 * x:2 * x + 2, c].ravel()
                s = np.sort(patch)[::-1]
                print("      at", (b, y, x, c), "ours", idx[b, y, x, c], "ref", ref_idx[b, y, x, c], "top two", s[0], s[1], "gap %.2e" % ((s[0] - s[1]) / max(abs(s[0]), 1e-30)))
