"""Accuracy of the forward conv implementations against a float64 reference: FFMA fp32 vs the split-operand
tcgen05 path (DCV_FWD_TC=1).  Prints rel-L2 and the mean SIGNED relative error (a bias shows truncating accumulation)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.nn.functional as F
from deconv_api_b200 import engine as E

rng = np.random.default_rng(0)
for (hw, cin, cout) in ((14, 512, 512), (56, 128, 256), (112, 64, 64)):
    x = np.maximum(rng.normal(0.5, 1, (1, hw, hw, 16)), 0).astype(np.float32)
    w0 = rng.normal(0, 0.1, (3, 3, 16, cin)).astype(np.float32)
    W = (rng.normal(0, 1, (3, 3, cin, cout)) * np.sqrt(2.0 / (9 * cin))).astype(np.float32)
    b = rng.normal(0, 0.1, (cout,)).astype(np.float32)
    eng = E.Engine(0)
    eng.set_forward_mode(os.environ.get("DCV_FWD_TC", "1") == "1")
    eng.add_conv("c0", w0, np.zeros(cin, np.float32), relu=True)     # cin = 16 -> also on the tc path
    eng.add_conv("c1", W, b, relu=False)
    eng.forward(x, 1)
    a0 = eng.activation(0)
    y = eng.activation(1)
    ref = F.conv2d(torch.from_numpy(a0).double().permute(0, 3, 1, 2), torch.from_numpy(W).double().permute(3, 2, 0, 1),
                   torch.from_numpy(b).double(), padding=1).permute(0, 2, 3, 1).numpy()
    err = (y.astype(np.float64) - ref)
    big = np.abs(ref) > 0.1 * np.abs(ref).mean()
    print("hw %3d %3d->%3d  rel-L2 %.2e   mean signed rel err %+.2e   (DCV_FWD_TC=%s)" % (
        hw, cin, cout, np.linalg.norm(err) / np.linalg.norm(ref), np.mean(err[big] / np.abs(ref[big])), os.environ.get("DCV_FWD_TC", "1")))
    eng.close()
