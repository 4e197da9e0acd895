"""How much do the per-launch CUDA event pairs (dcv_profile) cost the 512-projection step?  Same step, events on / off."""
import os, sys, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from deconv_api_b200 import synth, vgg16
model = vgg16.VGG16(weights=None, include_top=False, input_shape=(224, 224, 3))
eng = model.bound_engine(0, upto="block4_conv3"); li = eng.index["block4_conv3"]
img = torch.from_numpy(synth.image_batch(1)).cuda()
feats = np.arange(512, dtype=np.int32)
out8 = torch.empty(512 * 224 * 224 * 3, dtype=torch.uint8, device="cuda")
out32 = torch.empty(512 * 224 * 224 * 3, dtype=torch.float32, device="cuda")
stream = torch.cuda.ExternalStream(eng.stream, device=0)
def run(fmt, prof, steps=5):
    out = out8 if fmt == "u8" else out32
    def step():
        eng.forward_dev(img.data_ptr(), 1, 224, 224, li)
        eng.project_dev(li, feats, out.data_ptr(), precision="bf16", out_format=fmt)
    for _ in range(3): step()
    eng.sync(); eng.profile(prof)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps): step()
    e1.record(stream); eng.sync(); eng.profile(False)
    return e0.elapsed_time(e1) / steps
for rep in range(2):
    for fmt in ("u8", "f32"):
        for prof in (True, False):
            print("%-3s events %-5s %.3f ms/step" % (fmt, prof, run(fmt, prof)))
