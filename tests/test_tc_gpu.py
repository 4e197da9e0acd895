"""GPU parity of the tensor-core (tcgen05) backward chain, DCV_BF16: bf16 operands, fp32 accumulation.
Tolerances: against the oracle evaluated on bf16-ROUNDED operands the kernel must agree to fp32
accumulation noise (1e-5: this pins the implicit-GEMM indexing, swizzle and halo handling exactly);
against the unrounded fp32 oracle north_star's bf16 tolerance applies (rel-L2 <= 2e-2)."""
import numpy as np
import pytest
import torch

from oracle import keras_shim, restatement as rs
from deconv_api_b200 import deepdream as dd, synth, vgg16
from conftest import rel_l2

pytestmark = pytest.mark.gpu

BF16_TOL = 2e-2


def bf16_round(a):
    return torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(torch.bfloat16).to(torch.float32).numpy()


@pytest.mark.parametrize("B,H,W,cin,cout,relu", [
    (1, 8, 8, 64, 64, True),               # one k-block per tap, N tile 64
    (2, 14, 14, 128, 256, True),           # multi k-block, items sharing tiles, N tile 128
    (1, 28, 28, 512, 512, True),           # block4 shape: 8 k-blocks x 9 taps, two N tiles of 256
    (3, 10, 6, 256, 64, False),            # N tile 256, linear
    (1, 16, 16, 3, 64, True),              # last backward layer: 64 -> 3 channels (N tile 16, fp32 compact out)
    (2, 9, 7, 5, 24, True),                # ragged channels and odd sizes: padded to 64
    (1, 224, 224, 3, 64, True),
    (300, 4, 4, 64, 64, True),             # many tiny items: tiles straddle many item boundaries
    # large enough (>= 74 tiles of 256 rows) for the CTA-pair kernels (tc_kernel2.cuh), one per staging mode:
    (40, 28, 28, 256, 512, True),          # pair, tap mode, N tile 256
    (64, 28, 28, 128, 256, True),          # pair, band mode, N tile 128
    (24, 56, 56, 64, 128, False),          # pair, band + resident half-weights, N tile 64, K = 128
    (8, 112, 112, 64, 64, True),           # pair, band + resident, K = 64 (block1_conv2 shape)
])
def test_deconv3x3_bf16_vs_oracle(engine, B, H, W, cin, cout, relu):
    rng = np.random.default_rng(B * 1000 + H + cin + cout + 7)
    x = np.maximum(rng.normal(0, 1, (B, H, W, cout)), 0).astype(np.float32)
    lim = np.sqrt(6.0 / (9 * cin + 9 * cout))
    Wt = rng.uniform(-lim, lim, (3, 3, cin, cout)).astype(np.float32)
    out = engine.deconv3x3(x, Wt, relu=relu, precision="bf16")
    assert out.shape == (B, H, W, cin)
    exact = rs.conv_down(bf16_round(x), bf16_round(Wt), relu=relu)
    assert rel_l2(out, exact) < 1e-5
    assert rel_l2(out, rs.conv_down(x, Wt, relu=relu)) < BF16_TOL
    if relu:
        assert out.min() >= 0


def test_mini_model_bf16(golden):
    g = golden("mini_reference.npz")
    mini = keras_shim.mini_model(seed=7)
    for tgt in ("predictions", "block2_pool", "block1_conv2"):
        res = dd.visualize_all_layers(mini, g["x"].copy(), tgt, "all", precision="bf16")
        for name, lst in res.items():
            ref = g["all|%s|%s" % (tgt, name)]
            assert len(lst) == len(ref)
            for a, b in zip(lst, ref):
                assert a.shape == b.shape and rel_l2(a, b) < BF16_TOL, (tgt, name)


@pytest.fixture(scope="module")
def vgg():
    return vgg16.VGG16(weights=None, include_top=True), synth.image_batch(1)


def test_vgg16_block5_conv3_bf16_vs_reference_golden(vgg, golden):
    g = golden("vgg16_block5_conv3_reference.npz")
    model, img = vgg
    layers = ["block5_conv3", "block4_pool", "block3_conv3", "block1_conv2", "block1_conv1"]
    res = dd.visualize_all_layers(model, img, "block5_conv3", "all", precision="bf16", only_layers=layers)
    worst = 0.0
    for key in g.files:
        if key.startswith("proj|"):
            _, name, k = key.split("|")
            a = res[name][int(k)]
            assert a.shape == (224, 224, 3) and a.min() >= 0
            worst = max(worst, rel_l2(a, g[key]))
    assert worst < BF16_TOL, worst
    # the forward tape stays fp32 in both modes, so the switches are those of the fp32 test
    for name in layers:
        st = np.array([[a.astype(np.float64).sum(), (a.astype(np.float64) ** 2).sum()] for a in res[name]])
        assert np.allclose(st, g["stats|%s" % name][:, :2], rtol=5e-2), name


def test_config2_slice_bf16_vs_fp32(vgg):
    """BASELINE configs[1] on a slice: 40 feature maps of block4_conv3, bf16 chain vs fp32 chain vs oracle."""
    model, img = vgg
    eng = model.bound_engine(upto="block4_conv3")
    li = eng.index["block4_conv3"]
    eng.forward(img, li)
    sums = eng.channel_sums(li)
    feats = np.argsort(-sums, kind="stable")[:40].astype(np.int32)      # the 40 strongest maps, strongest first
    a = eng.project(li, feats, precision="bf16")
    b = eng.project(li, feats, precision="fp32")
    assert a.shape == b.shape == (40, 1, 224, 224, 3)
    rels = [rel_l2(a[k], b[k]) for k in range(40) if np.abs(b[k]).max() > 0]
    assert max(rels) < BF16_TOL, max(rels)
    ref = rs.visualize_all_layers(keras_shim.VGG16(include_top=False, input_shape=(224, 224, 3)), img, "block4_conv3",
                                  "all", top=4, only_layers=["block4_conv3"])["block4_conv3"]
    for k in range(4):
        assert rel_l2(b[k, 0], ref[k]) < 1e-5 and rel_l2(a[k, 0], ref[k]) < BF16_TOL
    # linearity in the seed does not hold (ReLU), but idempotence of the call does: same inputs, same bits
    a2 = eng.project(li, feats, precision="bf16")
    assert np.array_equal(a, a2)


def test_mini_model_batch2_bf16(golden):
    g = golden("mini_reference.npz")
    mini = keras_shim.mini_model(seed=7)
    res = dd.visualize_all_layers(mini, g["x2"].copy(), "block2_conv1", "all", precision="bf16")
    for name, lst in res.items():
        ref = g["all|b2|%s" % name]
        assert np.stack(lst).shape == ref.shape
        assert rel_l2(np.stack(lst), ref) < BF16_TOL, name


def test_small_workspace_chunks_give_identical_bits(vgg, monkeypatch):
    """A tiny workspace budget forces the projections through many chunks (and, for host output, through the
    pipelined copy path); the bytes must be identical to the single-chunk result."""
    from deconv_api_b200 import engine as E
    model, img = vgg
    eng = model.bound_engine(upto="block3_conv3")
    li = eng.index["block3_conv3"]
    eng.forward(img, li)
    feats = np.arange(0, 96, dtype=np.int32)
    want = {p: eng.project(li, feats, precision=p) for p in ("bf16", "fp32")}
    monkeypatch.setenv("DCV_WORKSPACE_MB", "256")
    small = E.Engine(0)
    try:
        for l in model.layers[1:]:
            if l.kind == "conv":
                W, b = l.get_weights()
                small.add_conv(l.name, W, b, relu=True)
            elif l.kind == "pool":
                small.add_pool(l.name)
            if l.name == "block3_conv3":
                break
        small.forward(img, small.index["block3_conv3"])
        for p in ("bf16", "fp32"):
            got = small.project(small.index["block3_conv3"], feats, precision=p)
            assert np.array_equal(got, want[p]), p
    finally:
        small.close()


def test_projection_non_square_image_vs_oracle():
    """A 96x160 image (sizes the golden files do not cover), 24 maps of block3_conv3 with bias-carrying random weights:
    the 64-channel layers run the dx-grouped kernel with its TMA-store epilogue (ragged last tile: 24*97*161 rows are
    not a multiple of 120), the fp32 mode runs the split-operand kernels incl. the exact GEMM + col2im last layer.

    Switches: this input has one pooling window whose two largest values differ by 7e-7 relative (3 ulp); two correct
    fp32 convolutions with different summation orders (ours, torch's, the reference's Eigen) may rank them differently.
    The test therefore requires every switch that differs from the oracle's to be such a near-tie (gap <= 1e-5), and
    then lets the oracle follow the library's choice so that the projection chain itself is held to 1e-5."""
    model = vgg16.VGG16(weights=None, include_top=False, input_shape=(96, 160, 3), seed=5, bias=True)
    ref = keras_shim.VGG16(include_top=False, input_shape=(96, 160, 3), seed=5, bias=True)
    img = synth.caffe_preprocess(synth.smooth_image_u8(3, 96, 160))
    eng = model.bound_engine(upto="block3_conv3")
    li = eng.index["block3_conv3"]
    eng.forward(img, li)
    tape = rs.FastTape(ref, "block3_conv3", img)
    flips = 0
    for name, op_i in zip(("block1_pool", "block2_pool"), [i for i, op in enumerate(tape.ops) if op[0] == "pool"]):
        _, idx, (H, W) = tape.ops[op_i]
        flat = idx[0].numpy()                                            # (C, h, w): index into the H x W plane
        sw_ref = (((flat // W) % 2) * 2 + (flat % W) % 2).transpose(1, 2, 0)
        sw = eng.switch_index(eng.index[name])[0].astype(np.int64)       # (h, w, C)
        act = eng.activation(eng.index[name.replace("pool", "conv2")])[0]
        for (y, x, c) in np.argwhere(sw != sw_ref):
            win = act[2 * y:2 * y + 2, 2 * x:2 * x + 2, c].ravel()
            assert abs(win[sw[y, x, c]] - win[sw_ref[y, x, c]]) <= 1e-5 * np.abs(win).max(), (name, y, x, c, win)
            flips += 1
        h, w, C = sw.shape
        yy, xx = np.meshgrid(np.arange(h), np.arange(w), indexing="ij")
        follow = ((2 * yy[..., None] + sw // 2) * W + 2 * xx[..., None] + sw % 2).transpose(2, 0, 1)[None]
        tape.ops[op_i] = ("pool", torch.from_numpy(np.ascontiguousarray(follow)), (H, W))
    assert flips <= 4
    sums = eng.channel_sums(li)
    feats = np.argsort(-sums, kind="stable")[:24].astype(np.int32)
    want = tape.project([int(f) for f in feats])
    got32 = eng.project(li, feats, precision="fp32")[:, 0]
    got16 = eng.project(li, feats, precision="bf16")[:, 0]
    assert got32.shape == want.shape == (24, 96, 160, 3)
    for k in range(24):
        if np.abs(want[k]).max() > 0:
            assert rel_l2(got32[k], want[k]) < 1e-5, k
            assert rel_l2(got16[k], want[k]) < BF16_TOL, k
