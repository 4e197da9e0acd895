"""Generates the golden fixtures in this directory by running the REFERENCE's own files
(/root/reference/app/deepdream.py, unmodified) on the Keras shim of oracle/keras_shim.

Run in the build container only (the reference does not exist on the GPU box):
    python tests/golden/make_golden.py
The reference has no tests or golden vectors of its own; these files are the pins.
"""
import contextlib
import io
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
HERE = os.path.dirname(os.path.abspath(__file__))

from oracle import keras_shim, reference_loader  # noqa: E402
from deconv_api_b200 import synth  # noqa: E402


def quiet(fn, *a, **k):
    with contextlib.redirect_stdout(io.StringIO()):
        return fn(*a, **k)


def pack2(idx):
    """(…) uint8 values 0..3 -> packed bytes, 4 per byte (little end first)."""
    flat = idx.reshape(-1).astype(np.uint8)
    pad = (-flat.size) % 4
    flat = np.concatenate([flat, np.zeros(pad, np.uint8)])
    q = flat.reshape(-1, 4)
    return (q[:, 0] | (q[:, 1] << 2) | (q[:, 2] << 4) | (q[:, 3] << 6)).astype(np.uint8)


def switch_to_idx(sw):
    """reference one-hot float64 mask (B,H,W,C) -> (B,H/2,W/2,C) index r*2+c."""
    B, H, W, C = sw.shape
    oh, ow = H // 2, W // 2
    idx = np.zeros((B, oh, ow, C), np.uint8)
    cnt = np.zeros((B, oh, ow, C), np.int32)
    for k in range(4):
        r, q = divmod(k, 2)
        m = sw[:, r:2 * oh:2, q:2 * ow:2, :] != 0
        idx[m] = k
        cnt += m
    assert (cnt == 1).all(), "reference switch is not one-hot per patch"
    return idx


def main():
    dd = reference_loader.load_deepdream()

    # ---- 1. tiny model, every target incl. dense/flatten, both modes: complete outputs -------------
    mini = keras_shim.mini_model(seed=7)
    x = np.random.default_rng(0).normal(0, 50, (1, 16, 16, 3)).astype(np.float32)
    out = {"x": x}
    for mode in ("all", "max"):
        for tgt in ("predictions", "block2_pool", "block1_conv2"):
            res = quiet(dd.visualize_all_layers, mini, x.copy(), tgt, mode)
            for name, lst in res.items():
                out["%s|%s|%s" % (mode, tgt, name)] = np.stack(lst) if lst else np.zeros((0,), np.float32)
    # batch of 2 (the function supports it; squeeze keeps the batch axis)
    x2 = np.random.default_rng(1).normal(0, 50, (2, 16, 16, 3)).astype(np.float32)
    out["x2"] = x2
    res = quiet(dd.visualize_all_layers, mini, x2.copy(), "block2_conv1", "all")
    for name, lst in res.items():
        out["all|b2|%s" % name] = np.stack(lst)
    np.savez_compressed(os.path.join(HERE, "mini_reference.npz"), **out)

    # ---- 2. DPooling unit vectors straight from the reference class ------------------------------
    class _L(object):
        pool_size = (2, 2)
    rng = np.random.default_rng(5)
    xin = np.maximum(rng.normal(0, 1, (2, 12, 10, 6)), 0).astype(np.float32)      # ReLU'd: many all-zero patches
    xin = np.round(xin * 4) / 4                                                  # quantised: many ties
    xin[0, 0:2, 0:2, 0] = [[0, 5], [5, 5]]                                        # KAT: first max in row-major -> (0,1)
    xin[0, 2:4, 0:2, 0] = 0                                                       # KAT: all-zero patch -> (0,0)
    xin[0, 4:6, 0:2, 0] = [[1, 1], [7, 7]]                                        # KAT: tie in the second row -> (1,0)
    p = dd.DPooling(_L())
    pooled = p.up(xin)
    g = rng.normal(0, 1, pooled.shape).astype(np.float32)
    unpooled = p.down(g)
    np.savez_compressed(os.path.join(HERE, "pool_reference.npz"), x=xin, pooled=pooled, switch=p.switch.astype(np.uint8),
                        idx=switch_to_idx(p.switch), g=g, unpooled=unpooled)

    # ---- 3. VGG16 @224, block5_conv3: top-8 lists for all 17 layers, switches, a few projections ---
    from keras.applications import vgg16
    model = vgg16.VGG16(weights="imagenet", include_top=True)
    img = synth.image_batch(1)
    # capture the wrappers' state: monkey-patch nothing — rerun the forward with the reference classes
    res = quiet(dd.visualize_all_layers, model, img.copy(), "block5_conv3", "all")
    vg = {}
    layer_order = list(res.keys())
    vg["layers"] = np.array(layer_order)
    vg["counts"] = np.array([len(res[k]) for k in layer_order])
    keep = {"block5_conv3": [0, 1], "block4_pool": [0], "block3_conv3": [0], "block1_conv2": [0], "block1_conv1": [7]}
    for name, ks in keep.items():
        for k in ks:
            vg["proj|%s|%d" % (name, k)] = res[name][k]
    # per-projection checksums for everything else (sum, sum of squares, max) in float64
    for name in layer_order:
        st = np.array([[a.astype(np.float64).sum(), (a.astype(np.float64) ** 2).sum(), a.max()] for a in res[name]])
        vg["stats|%s" % name] = st
    # forward-side pins from the reference classes themselves
    x = img
    tops = {}
    for layer in model.layers[1:]:
        if layer.name == "block5_pool":
            break
        if type(layer).__name__ == "Conv2D":
            w = dd.DConvolution2D(layer)
            x = w.up(x)
        else:
            w = dd.DPooling(layer)
            x = w.up(x)
            vg["switch|%s" % layer.name] = pack2(switch_to_idx(w.switch))
        tf = dd.find_top_filters(x)
        tops[layer.name] = tf
        vg["top_idx|%s" % layer.name] = np.array([t[0] for t in tf], np.int32)
        vg["top_sum|%s" % layer.name] = np.array([t[1] for t in tf], np.float64)
        vg["act_stats|%s" % layer.name] = np.array([np.asarray(x, np.float64).sum(), (np.asarray(x, np.float64) ** 2).sum()])
    np.savez_compressed(os.path.join(HERE, "vgg16_block5_conv3_reference.npz"), **vg)

    # ---- 4. deprocess_image on the API's 2x2 mosaic (main.py:67-72) and on a small array ------------
    l = res["block5_conv3"]
    mosaic = np.concatenate((np.concatenate((l[0], l[1]), axis=1), np.concatenate((l[2], l[3]), axis=1)), axis=0)
    m_in = mosaic.copy()
    m_out = dd.deprocess_image(mosaic)
    small = np.random.default_rng(3).normal(0.3, 2.0, (37, 29, 3)).astype(np.float32)
    s_in = small.copy()
    s_out = dd.deprocess_image(small)
    # crop of the mosaic input (full one is 2.4 MB): the test recomputes the mosaic from its own projections,
    # so only the outputs and the mutated argument are stored
    np.savez_compressed(os.path.join(HERE, "deprocess_reference.npz"), mosaic_u8=m_out, mosaic_in_stats=np.array(
        [m_in.astype(np.float64).mean(), m_in.astype(np.float64).std()]), small_in=s_in, small_u8=s_out, small_mutated=small)
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print("%-44s %8.1f KB" % (f, os.path.getsize(os.path.join(HERE, f)) / 1024.0))


if __name__ == "__main__":
    main()
