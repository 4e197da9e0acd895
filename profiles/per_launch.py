"""Per-launch device times of one config-2 step (CUDA events inside the library), for profiles/*.md.
    python profiles/per_launch.py [bf16|fp32] [nfeat]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from deconv_api_b200 import synth, vgg16

prec = sys.argv[1] if len(sys.argv) > 1 else "bf16"
nfeat = int(sys.argv[2]) if len(sys.argv) > 2 else 512
model = vgg16.VGG16(weights=None, include_top=False, input_shape=(224, 224, 3))
eng = model.bound_engine(0, upto="block4_conv3")
li = eng.index["block4_conv3"]
img = synth.image_batch(1)
feats = np.arange(nfeat, dtype=np.int32)
import torch
out = torch.empty(nfeat * 224 * 224 * 3, dtype=torch.float32, device="cuda")
for _ in range(3):
    eng.forward(img, li); eng.project_dev(li, feats, out.data_ptr(), precision=prec)
eng.sync()
eng.profile(True)
eng.forward(img, li); eng.project_dev(li, feats, out.data_ptr(), precision=prec)
rows = eng.profile_launches()
eng.profile(False)
tot = sum(r[1] for r in rows)
print("%-3s %-10s %9s %10s %10s %6s" % ("#", "class", "ms", "TFLOP/s", "GB/s", "share"))
for i, (c, ms, fl, by) in enumerate(rows):
    print("%-3d %-10s %9.4f %10.1f %10.1f %5.1f%%" % (i, c, ms, fl / ms / 1e9 if ms else 0, by / ms / 1e6 if ms else 0, 100 * ms / tot))
print("total %.3f ms over %d launches" % (tot, len(rows)))
