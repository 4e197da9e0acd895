"""GPU parity of the DeepDream row (a12).  No reference implementation exists: the oracle is
oracle/dream_oracle.py (torch autograd on the Appendix-A spec) — parity against the reference is unpinned."""
import numpy as np
import pytest

from oracle import dream_oracle as do, keras_shim
from deconv_api_b200 import dream as dm, engine as E, synth, vgg16
from conftest import rel_l2

pytestmark = pytest.mark.gpu


def test_resize_and_roll_vs_oracle(engine):
    rng = np.random.default_rng(0)
    for (h, w), (oh, ow) in (((37, 53), (61, 44)), ((64, 64), (45, 45)), ((365, 365), (512, 512)), ((20, 30), (20, 30))):
        x = rng.normal(0, 50, (2, h, w, 3)).astype(np.float32)
        got = engine.resize_bilinear(x, (oh, ow))
        assert rel_l2(got, do.resize(x, (oh, ow))) < 1e-6
    x = rng.normal(0, 1, (2, 17, 23, 3)).astype(np.float32)
    for oy, ox in ((0, 0), (5, -7), (-32, 32), (17, 23), (40, -50)):
        assert np.array_equal(engine.roll(x, oy, ox), np.roll(x, (oy, ox), axis=(1, 2)))


@pytest.mark.parametrize("hw", [(64, 64), (45, 61)])          # odd sizes: pools floor, trailing row/col get zero gradient
def test_gradient_vs_autograd(hw):
    model = vgg16.VGG16(weights=None, include_top=False, input_shape=(None, None, 3), seed=9, bias=True)
    ref = keras_shim.VGG16(include_top=False, input_shape=(None, None, 3), seed=9, bias=True)
    layers = {"block2_conv2": 0.7, "block3_pool": 1.0, "block4_conv2": 2.0}
    img = np.concatenate([synth.caffe_preprocess(synth.smooth_image_u8(s, *hw)) for s in (1, 2)], axis=0)
    eng = model.bound_engine(upto="block4_conv2")
    g, loss = eng.dream_gradient(img, [eng.index[n] for n in layers], list(layers.values()))
    loss_ref, g_ref = do.loss_and_grad(do._convs(ref, "block4_conv2"), img, layers)
    assert abs(loss - loss_ref) <= 1e-5 * abs(loss_ref)
    assert rel_l2(g, g_ref) < 1e-5


def test_dream_loop_vs_oracle():
    model = vgg16.VGG16(weights=None, include_top=False, input_shape=(None, None, 3), seed=9)
    ref = keras_shim.VGG16(include_top=False, input_shape=(None, None, 3), seed=9)
    layers = {"block2_conv2": 0.5, "block3_conv3": 2.0}
    img = synth.caffe_preprocess(synth.smooth_image_u8(4, 96, 80))
    out, losses = dm.dream(model, img, layers=layers, num_octave=3, octave_scale=1.4, iterations=4, step=1.0, jitter=8,
                           seed=3, return_losses=True)
    shapes = dm.octave_shapes(96, 80, 3, 1.4)
    assert shapes == [(48, 40), (68, 57), (96, 80)]
    jt = dm.jitter_offsets(12, 8, 3)
    out_ref, losses_ref = do.dream(ref, img, layers, shapes, 4, 1.0, jt)
    assert out.shape == img.shape
    assert np.allclose(losses, losses_ref, rtol=1e-4)
    delta, delta_ref = out - img, out_ref - img
    assert np.linalg.norm(delta_ref) > 1.0                      # the loop really changed the image
    assert rel_l2(out, out_ref) < 1e-5
    assert rel_l2(delta, delta_ref) < 1e-3                      # gradient ascent amplifies fp32 noise; stated tolerance
    # single (H,W,3) image in, same shape out
    o1 = dm.dream(model, img[0], layers=layers, num_octave=2, iterations=1, jitter=0)
    assert o1.shape == img.shape[1:]


def test_gradient_bf16_tensor_core_path():
    """Forward and backward on tcgen05 (bf16 tape).  Against the fp32 autograd oracle: north_star's bf16 tolerance
    does not apply to a raw gradient (ReLU masks / switches flip with bf16 activations), so the stated bound here is
    rel-L2 <= 0.15 and cosine >= 0.99; the dream IMAGE is held to 2e-2 below."""
    model = vgg16.VGG16(weights=None, include_top=False, input_shape=(None, None, 3), seed=9, bias=True)
    ref = keras_shim.VGG16(include_top=False, input_shape=(None, None, 3), seed=9, bias=True)
    layers = {"block2_conv2": 0.7, "block3_pool": 1.0, "block4_conv2": 2.0}
    for hw in ((64, 64), (45, 61)):
        img = np.concatenate([synth.caffe_preprocess(synth.smooth_image_u8(s, *hw)) for s in (1, 2)], axis=0)
        eng = model.bound_engine(upto="block4_conv2")
        g, loss = eng.dream_gradient(img, [eng.index[n] for n in layers], list(layers.values()), precision="bf16")
        loss_ref, g_ref = do.loss_and_grad(do._convs(ref, "block4_conv2"), img, layers)
        assert abs(loss - loss_ref) <= 2e-2 * abs(loss_ref)
        cos = float((g * g_ref).sum() / (np.linalg.norm(g) * np.linalg.norm(g_ref)))
        assert cos > 0.99 and rel_l2(g, g_ref) < 0.15, (hw, cos, rel_l2(g, g_ref))


def test_dream_loop_bf16_vs_oracle():
    model = vgg16.VGG16(weights=None, include_top=False, input_shape=(None, None, 3), seed=9)
    ref = keras_shim.VGG16(include_top=False, input_shape=(None, None, 3), seed=9)
    layers = {"block2_conv2": 0.5, "block3_conv3": 2.0}
    img = synth.caffe_preprocess(synth.smooth_image_u8(4, 96, 80))
    out = dm.dream(model, img, layers=layers, num_octave=3, iterations=4, step=1.0, jitter=8, seed=3, precision="bf16")
    out_ref, _ = do.dream(ref, img, layers, dm.octave_shapes(96, 80, 3, 1.4), 4, 1.0, dm.jitter_offsets(12, 8, 3))
    assert rel_l2(out, out_ref) < 2e-2                      # north_star: dream images within 2e-2 in bf16
    assert rel_l2(out - img, out_ref - img) < 0.2           # and the applied change itself agrees to 20 %


def test_gradient_bf16_large_image_kernel_variants():
    """At >= 134x134 pixels the 64-channel layers switch to the dx-grouped kernel (bias epilogue in the forward,
    prep epilogue in the backward, output through a shared-memory tile + TMA store) and block2+ use multi-wave grids;
    the first layer runs as im2col + GEMM.  Odd size on purpose (ragged last tile, pools floor)."""
    model = vgg16.VGG16(weights=None, include_top=False, input_shape=(None, None, 3), seed=9, bias=True)
    ref = keras_shim.VGG16(include_top=False, input_shape=(None, None, 3), seed=9, bias=True)
    layers = {"block1_conv2": 0.3, "block2_conv2": 0.7, "block3_conv2": 1.5}
    hw = (171, 150)
    img = synth.caffe_preprocess(synth.smooth_image_u8(7, *hw))
    eng = model.bound_engine(upto="block3_conv2")
    g, loss = eng.dream_gradient(img, [eng.index[n] for n in layers], list(layers.values()), precision="bf16")
    loss_ref, g_ref = do.loss_and_grad(do._convs(ref, "block3_conv2"), img, layers)
    assert abs(loss - loss_ref) <= 2e-2 * abs(loss_ref)
    cos = float((g * g_ref).sum() / (np.linalg.norm(g) * np.linalg.norm(g_ref)))
    assert cos > 0.99 and rel_l2(g, g_ref) < 0.15, (cos, rel_l2(g, g_ref))
    # and the fp32 mode (tensor cores, split operands) at the same size against the same oracle
    g32, loss32 = eng.dream_gradient(img, [eng.index[n] for n in layers], list(layers.values()), precision="fp32")
    assert abs(loss32 - loss_ref) <= 1e-5 * abs(loss_ref)
    assert rel_l2(g32, g_ref) < 1e-5


def test_dream_after_set_weights_does_not_replay_a_stale_graph():
    """A cached CUDA graph of the bf16 dream holds raw pointers into the tape and the packed weights.  Changing the
    weights makes the model rebuild its engine network (dcv_net_clear frees those buffers): the next dream with the
    SAME key (shape, losses, step) must re-capture, and equal a fresh model that was given the new weights up front."""
    layers = {"block2_conv2": 0.5, "block3_conv3": 2.0}
    img = synth.caffe_preprocess(synth.smooth_image_u8(4, 96, 80))
    kw = dict(layers=layers, num_octave=2, iterations=4, step=1.0, jitter=8, seed=3, precision="bf16")
    model = vgg16.VGG16(weights=None, include_top=False, input_shape=(None, None, 3), seed=9)
    first = dm.dream(model, img, **kw)
    other = vgg16.VGG16(weights=None, include_top=False, input_shape=(None, None, 3), seed=21)
    for dst, src in zip(model.layers, other.layers):
        if dst.kind == "conv":
            dst.set_weights(src.get_weights())
    second = dm.dream(model, img, **kw)                  # same engine, network cleared and re-added
    fresh = dm.dream(vgg16.VGG16(weights=None, include_top=False, input_shape=(None, None, 3), seed=21), img, **kw)
    assert rel_l2(second, fresh) < 1e-6
    assert rel_l2(second - img, first - img) > 0.1       # the new weights really changed the result


def test_gradient_at_config3_size_512():
    """BASELINE configs[2] runs 512 x 512 dreams on the mixed-layer loss block3_conv3 / block4_conv3 / block5_conv3: one
    gradient at exactly that size and loss against torch autograd (a few seconds of CPU), both precisions.  Of the 7.8
    million pooling windows of this image 8 are near-ties that the library's and torch's fp32 convs rank differently
    (each moves a gradient path: 1e-3 in rel-L2, profiles/diag_dream512.py); as in the projection tests every differing
    switch must be a near-tie (<= 1e-5 of the layer's largest activation) and the oracle then routes those windows
    like the library.  2048 x 2048 (configs[4]) is the same code on 16 x the pixels; its autograd oracle needs minutes
    of CPU and is not run here."""
    import torch
    import torch.nn.functional as F
    model = vgg16.VGG16(weights=None, include_top=False, input_shape=(None, None, 3))
    ref = keras_shim.VGG16(include_top=False, input_shape=(None, None, 3))
    layers = dict(dm.DEFAULT_LAYERS)
    img = synth.caffe_preprocess(synth.smooth_image_u8(0, 512, 512))
    eng = model.bound_engine(upto="block5_conv3")
    idx, coeffs = [eng.index[n] for n in layers], list(layers.values())
    g32, loss32 = eng.dream_gradient(img, idx, coeffs, precision="fp32")
    ops = do._convs(ref, "block5_conv3")
    # the library's switches where they differ from torch's, each checked to be a near-tie
    follow, flips = {}, 0
    t = torch.from_numpy(img).permute(0, 3, 1, 2)
    for kind, name, W, b in ops:
        if kind == "conv":
            t = F.relu(F.conv2d(t, W, b, padding=1))
            continue
        wd = t.shape[3]
        act = t[0].permute(1, 2, 0).numpy()
        t, ind = F.max_pool2d(t, 2, 2, return_indices=True)
        flat = ind[0].numpy()
        sw_ref = (((flat // wd) % 2) * 2 + (flat % wd) % 2).transpose(1, 2, 0)
        sw = eng.switch_index(eng.index[name])[0].astype(np.int64)
        scale = float(np.abs(act).max())
        for (y, x, c) in np.argwhere(sw != sw_ref):
            win = act[2 * y:2 * y + 2, 2 * x:2 * x + 2, c].ravel()
            assert abs(win[sw[y, x, c]] - win[sw_ref[y, x, c]]) <= 1e-5 * scale, (name, y, x, c, win)
            flips += 1
        follow[name] = sw[None]
    assert flips <= 32
    loss_ref, g_ref = do.loss_and_grad(ops, img, layers, pool_idx=follow)
    assert abs(loss32 - loss_ref) <= 1e-5 * abs(loss_ref)
    # Stated tolerance at this size: 2e-3 (measured 6.5e-4 with the pool routes followed, 1.1e-3 without).  The true
    # back-propagation also masks by the SIGN of every forward activation; among 40 million of them a few sit within an
    # ulp of zero and the two fp32 forward passes disagree on the mask there - each such element gates a whole gradient
    # path, exactly like a pooling near-tie.  Where no such coincidence occurs (64 x 64, 45 x 61, 171 x 150 above) the
    # same code path is held to 1e-5.
    assert rel_l2(g32, g_ref) < 2e-3
    g16, loss16 = eng.dream_gradient(img, idx, coeffs, precision="bf16")
    assert abs(loss16 - loss_ref) <= 2e-2 * abs(loss_ref)
    # bf16 tape AND bf16 backward through all 13 convs (the shallower losses of the tests above reach cosine 0.99):
    # measured cosine 0.979 at this depth; stated bound 0.97 / rel-L2 0.25.  The raw gradient is not what north_star
    # bounds (the dream image is, at 2e-2: test_dream_loop_bf16_vs_oracle); this pins the direction of ascent.
    cos = float((g16 * g_ref).sum() / (np.linalg.norm(g16) * np.linalg.norm(g_ref)))
    assert cos > 0.97 and rel_l2(g16, g_ref) < 0.25, (cos, rel_l2(g16, g_ref))


def test_dream_at_config5_size_runs_and_is_deterministic():
    """configs[4] size (2048 x 2048, four octaves): no oracle at this size (minutes of CPU autograd), so the properties
    that do not depend on one — the loop changes the image, stays finite, keeps the shape, and two runs give the same bits
    (fixed jitter stream, deterministic reductions, CUDA-graph replay included)."""
    model = vgg16.VGG16(weights=None, include_top=False, input_shape=(None, None, 3))
    img = synth.caffe_preprocess(synth.smooth_image_u8(0, 2048, 2048))
    kw = dict(num_octave=4, iterations=3, precision="bf16", seed=1)
    a = dm.dream(model, img, **kw)
    b = dm.dream(model, img, **kw)
    assert a.shape == img.shape and np.isfinite(a).all()
    assert np.array_equal(a, b)
    assert rel_l2(a, img) > 1e-3


_SPLITK_SCRIPT = r"""
import sys, numpy as np
from deconv_api_b200 import synth, vgg16
model = vgg16.VGG16(weights=None, include_top=False, input_shape=(None, None, 3), seed=9, bias=True)
layers = {"block3_conv3": 1.0, "block4_conv3": 1.5, "block5_conv3": 2.0}
out = {}
for hw in ((64, 64), (200, 184)):
    img = synth.caffe_preprocess(synth.smooth_image_u8(5, *hw))
    eng = model.bound_engine(upto="block5_conv3")
    g, loss = eng.dream_gradient(img, [eng.index[n] for n in layers], list(layers.values()), precision="bf16")
    g2, loss2 = eng.dream_gradient(img, [eng.index[n] for n in layers], list(layers.values()), precision="bf16")
    assert np.array_equal(g, g2) and loss == loss2          # the cluster reduction adds in a fixed order
    out["g%d" % hw[0]] = g
    out["l%d" % hw[0]] = np.float64(loss)
np.savez(sys.argv[1], **out)
"""


def test_cluster_split_k_variants_agree(tmp_path):
    """DeepDream's deep layers have fewer tiles than SMs; the band kernel then runs as clusters of 2 or 4 CTAs per tile
    that split K and reduce through CTA 0's shared memory (tc_kernel.cuh, SPLIT).  The rule picks the split per layer;
    DCV_TC_SPLITK = 2 / 4 forces it on every layer it divides (clusters then walk several tiles: the hand-back of the
    reduction buffer is exercised), 0 turns it off.  All variants must give the same gradient up to the fp32 summation
    order (outputs are bf16: a last-place flip here and there)."""
    import os, subprocess, sys
    res = {}
    for mode in ("0", "1", "2", "4"):
        path = str(tmp_path / ("g%s.npz" % mode))
        env = dict(os.environ, DCV_TC_SPLITK=mode, PYTHONPATH=os.pathsep.join(sys.path))
        subprocess.run([sys.executable, "-c", _SPLITK_SCRIPT, path], check=True, env=env, timeout=300)
        res[mode] = np.load(path)
    for mode in ("1", "2", "4"):
        for k in ("g64", "g200"):
            assert rel_l2(res[mode][k], res["0"][k]) < 1e-2, (mode, k, rel_l2(res[mode][k], res["0"][k]))
        for k in ("l64", "l200"):
            assert abs(res[mode][k] - res["0"][k]) <= 1e-3 * abs(res["0"][k])


def test_resize_kernel_variants_are_bit_equal():
    """The pipelined resize (bulk async copies into a shared-memory ring), the parity-slot, the strip and the
    one-thread-per-pixel kernels compute the same expression in the same order: equal bits on ten shapes (odd sizes,
    1 / 3 / 4 / 5 channels, enlargements and reductions up to 2.7x, 1 x 1 inputs, 2048 x 2048)."""
    import os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "profiles", "resize_variants.py")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "variants equal" in r.stdout, r.stdout[-2000:]
