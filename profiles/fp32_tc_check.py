"""fp32 projection mode: tensor-core (split-operand) backward convs against the FFMA kernels, all 136 projections of
the block5_conv3 sweep.  Usage: python profiles/fp32_tc_check.py  (spawns itself twice with DCV_FP32_TC=0 / 2)."""
import os, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

if len(sys.argv) > 1:
    from deconv_api_b200 import deepdream as dd, synth, vgg16
    model = vgg16.VGG16(weights=None, include_top=True)
    res = dd.visualize_all_layers(model, synth.image_batch(1), "block5_conv3", "all", precision="fp32")
    np.savez(sys.argv[1], **{"%s|%d" % (k, i): a for k, v in res.items() for i, a in enumerate(v)})
    sys.exit(0)

outs = {}
for mode in ("0", "1", "2"):
    path = "/tmp/fp32_tc_%s.npz" % mode
    subprocess.check_call([sys.executable, os.path.abspath(__file__), path], env=dict(os.environ, DCV_FP32_TC=mode),
                          stdout=subprocess.DEVNULL)
    outs[mode] = np.load(path)
ref = outs["0"]
for mode in ("1", "2"):
    worst = {}
    for key in ref.files:
        a, b = ref[key].astype(np.float64), outs[mode][key].astype(np.float64)
        e = np.sqrt(((a - b) ** 2).sum() / max((a ** 2).sum(), 1e-300))
        name = key.split("|")[0]
        worst[name] = max(worst.get(name, 0.0), e)
    print("DCV_FP32_TC=%s  max rel-L2 vs FFMA per layer:" % mode)
    for name, e in worst.items():
        print("   %-14s %.3e" % (name, e))
