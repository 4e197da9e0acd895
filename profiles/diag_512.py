"""Diagnostic: all 512 block4_conv3 projections (fp32 mode) against the oracle, worst maps, and the same maps projected alone."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from oracle import keras_shim
from deconv_api_b200 import synth, vgg16
from conftest import following_tape

def rel(a, b):
    return float(np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-30))

model = vgg16.VGG16(weights=None, include_top=False, input_shape=(224, 224, 3))
ref_model = keras_shim.VGG16(include_top=False, input_shape=(224, 224, 3))
img = synth.image_batch(1, 224, 224, seed0=0)
eng = model.bound_engine(upto="block4_conv3")
li = eng.index["block4_conv3"]
eng.forward(img, li)
tape, flips = following_tape(eng, ref_model, "block4_conv3", img)
print("flips", flips)
feats = np.arange(512, dtype=np.int32)
want = np.concatenate([tape.project([int(f) for f in feats[i:i + 32]]) for i in range(0, 512, 32)])
act = eng.activation(li)[0]
for prec in ("fp32", "bf16"):
    got = eng.project(li, feats, precision=prec)[:, 0]
    errs = np.array([rel(got[k], want[k]) if np.abs(want[k]).max() > 0 else 0.0 for k in range(512)])
    order = np.argsort(-errs)[:12]
    print(prec, "median %.2e  p99 %.2e  max %.2e" % (np.median(errs), np.percentile(errs, 99), errs.max()))
    for k in order:
        alone = eng.project(li, feats[k:k + 1], precision=prec)[0, 0]
        nz = int(np.count_nonzero(act[..., k]))
        print("   map %3d err %.2e | alone %.2e | alone vs batch %.2e | nonzero act %d, act max %.3e, proj max %.3e, proj nnz %d"
              % (k, errs[k], rel(alone, want[k]), rel(alone, got[k]), nz, act[..., k].max(), want[k].max(), np.count_nonzero(want[k])))
        if prec == "fp32" and k == order[0]:
            d = np.abs(got[k] - want[k])
            idx = np.unravel_index(np.argsort(-d.ravel())[:5], d.shape)
            for y, x, c in zip(*idx):
                print("        at", (y, x, c), "ours %.6e ref %.6e" % (got[k][y, x, c], want[k][y, x, c]))
