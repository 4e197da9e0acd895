"""Bit-equality of the resize kernel variants (DCV_RESIZE_PIXEL = 0 pipelined, 1 per pixel, 2 strip, 3 parity slots):
each variant runs in its own process (the switch is read once), the outputs are compared here."""
import os, subprocess, sys, tempfile
import numpy as np
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCRIPT = r"""
import sys, numpy as np
sys.path.insert(0, %r)
from deconv_api_b200 import engine as E
eng = E.Engine(0)
rng = np.random.default_rng(0)
out = {}
cases = [((2, 37, 53, 3), (52, 74)), ((1, 64, 64, 3), (23, 31)), ((3, 1, 1, 3), (5, 7)), ((1, 300, 17, 1), (111, 40)),
         ((1, 45, 61, 4), (45, 61)), ((2, 129, 257, 3), (181, 360)), ((1, 1463, 1463, 3), (2048, 2048)),
         ((1, 2048, 2048, 3), (747, 747)), ((1, 9, 2000, 3), (3, 5000)), ((1, 512, 512, 5), (366, 366))]
for i, (shp, ohw) in enumerate(cases):
    x = rng.standard_normal(shp).astype(np.float32)
    out["c%%d" %% i] = eng.resize_bilinear(x, ohw)
np.savez(sys.argv[1], **out)
""" % root
res = {}
with tempfile.TemporaryDirectory() as td:
    for mode in "0123":
        p = os.path.join(td, mode + ".npz")
        subprocess.run([sys.executable, "-c", SCRIPT, p], check=True, env=dict(os.environ, DCV_RESIZE_PIXEL=mode))
        res[mode] = dict(np.load(p))
bad = 0
for mode in "023":
    for k in res["1"]:
        same = np.array_equal(res[mode][k], res["1"][k])
        if not same:
            bad += 1
            print("mode", mode, k, "differs: max abs", float(np.abs(res[mode][k] - res["1"][k]).max()))
print("variants equal" if not bad else "%d differences" % bad)
