"""Diagnostic: projections of the dense head (predictions / fc2 / fc1) — library vs the fp32 oracle vs a float64 chain."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import keras_shim, restatement as rs, fp64_chain
from deconv_api_b200 import deepdream as dd, synth, vgg16

def rel(a, b):
    return float(np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-30))

model = vgg16.VGG16(weights=None, include_top=True)
ref_model = keras_shim.VGG16(include_top=True)
img = synth.image_batch(1)
layers = ["predictions", "fc2", "fc1", "flatten", "block5_pool"]
got = dd.visualize_all_layers(model, img, "predictions", "all", only_layers=layers, top=2, precision="fp32")
entries = rs.forward(rs.build_entries(ref_model, "predictions"), img)
names = [e.name for e in entries]
eng = model.bound_engine()
for lname in layers:
    i = names.index(lname)
    out = entries[i].up
    ours_act = eng.activation(eng.index[lname])
    print(lname, "forward activation rel-L2 vs oracle %.2e" % rel(ours_act.reshape(out.shape), out))
    for k, (f, _) in enumerate(rs.find_top_filters(out, 2)):
        seed = np.zeros_like(out)
        if out.ndim == 2: seed[:, f] = out[:, f]
        else: seed[..., f] = out[..., f]
        a32 = rs.backward(entries, i, seed)
        a64 = fp64_chain.backward64(entries, i, seed)
        print("   feat", f, "ours vs fp32 oracle %.2e | ours vs fp64 %.2e | fp32 oracle vs fp64 %.2e" % (rel(got[lname][k], a32), rel(got[lname][k], a64), rel(a32, a64)))
