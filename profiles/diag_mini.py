"""Diagnostic: mini model, every target of the golden file — library vs golden, per layer, worst projection."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import keras_shim, restatement as rs, fp64_chain
from deconv_api_b200 import deepdream as dd

def rel(a, b):
    return float(np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-30))

g = np.load(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "mini_reference.npz"))
mini = keras_shim.mini_model(seed=7)
x = g["x"].copy()
entries = rs.forward(rs.build_entries(mini, "predictions"), x)
names = [e.name for e in entries]
eng = dd.adopt(mini).bound_engine()
for fwd_tc in (True, False):
    eng.set_forward_mode(fwd_tc); eng.set_fp32_mode(fwd_tc)
    res = dd.visualize_all_layers(mini, x.copy(), "predictions", "all")
    print("tensor cores" if fwd_tc else "FFMA")
    for name, lst in res.items():
        ref = g["all|predictions|%s" % name]
        i = names.index(name)
        out = entries[i].up
        act = eng.activation(eng.index[name]).reshape(out.shape)
        errs = [rel(a, b) for a, b in zip(lst, ref)]
        top = rs.find_top_filters(out, 8)
        seed_err = [abs(float(act[..., f].sum()) - float(out[..., f].sum())) / abs(float(out[..., f].sum())) for f, _ in top]
        print("  %-14s worst %.2e  (forward act rel-L2 %.2e, worst seed-sum rel err %.2e)" % (name, max(errs), rel(act, out), max(seed_err)))
eng.set_forward_mode(True); eng.set_fp32_mode(True)
