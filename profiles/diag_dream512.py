"""Diagnostic: fp32 DeepDream gradient at 512x512 against torch autograd; switch agreement per pool layer."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.nn.functional as F
from oracle import dream_oracle as do, keras_shim
from deconv_api_b200 import dream as dm, synth, vgg16
def rel(a, b): return float(np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-30))
model = vgg16.VGG16(weights=None, include_top=False, input_shape=(None, None, 3))
ref = keras_shim.VGG16(include_top=False, input_shape=(None, None, 3))
layers = dict(dm.DEFAULT_LAYERS)
S = int(sys.argv[1]) if len(sys.argv) > 1 else 512
img = synth.caffe_preprocess(synth.smooth_image_u8(0, S, S))
eng = model.bound_engine(upto="block5_conv3")
idx, coeffs = [eng.index[n] for n in layers], list(layers.values())
loss_ref, g_ref = do.loss_and_grad(do._convs(ref, "block5_conv3"), img, layers)
g32, loss32 = eng.dream_gradient(img, idx, coeffs, precision="fp32")
print("scheme", os.environ.get("DCV_FP32_SCHEME", "1"), "loss rel", abs(loss32 - loss_ref) / abs(loss_ref), "grad rel-L2", rel(g32, g_ref))
d = np.abs(g32 - g_ref).sum(axis=-1)[0]
print("pixels with |diff| > 1e-3 * max|g|:", int((d > 1e-3 * np.abs(g_ref).max()).sum()), "of", d.size)
# switches of the library's tape against torch's
t = torch.from_numpy(img).permute(0, 3, 1, 2)
for kind, name, W, b in do._convs(ref, "block5_conv3"):
    if kind == "conv":
        t = F.relu(F.conv2d(t, W, b, padding=1))
    else:
        hw = t.shape[2:]
        t2, ind = F.max_pool2d(t, 2, 2, return_indices=True)
        Wd = hw[1]
        flat = ind[0].numpy()
        sw_ref = (((flat // Wd) % 2) * 2 + (flat % Wd) % 2).transpose(1, 2, 0)
        sw = eng.switch_index(eng.index[name])[0]
        bad = np.argwhere(sw != sw_ref)
        print(name, "switch mismatches", len(bad), "of", sw.size)
        t = t2
