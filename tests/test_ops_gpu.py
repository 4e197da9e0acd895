"""GPU parity of the individual operators, called through the C-ABI (ctypes) and checked against the
oracle restatement and the reference-generated golden vectors.  Bit-exact for switches, pooled values,
unpool placement; rel-L2 <= 1e-5 for fp32 conv arithmetic (north_star's stated tolerance)."""
import numpy as np
import pytest

from oracle import restatement as rs
from conftest import rel_l2

pytestmark = pytest.mark.gpu

FP32_TOL = 1e-5


def _relu_quant(rng, shape, q=4.0):
    x = np.maximum(rng.normal(0, 1, shape), 0).astype(np.float32)
    return np.round(x * q) / q


def test_pool_golden_bit_exact(engine, golden):
    g = golden("pool_reference.npz")
    pooled, idx = engine.maxpool_switch(g["x"])
    assert np.array_equal(pooled.astype(np.float64), g["pooled"])
    assert np.array_equal(idx, g["idx"])
    un = engine.unpool_switch(g["g"], idx)
    assert np.array_equal(un.astype(np.float64), g["unpooled"])


@pytest.mark.parametrize("shape", [(1, 224, 224, 64), (2, 28, 28, 512), (3, 14, 10, 6), (1, 8, 8, 3), (1, 2, 2, 1),
                                   (1, 9, 7, 5), (2, 16, 16, 130)])
def test_pool_unpool_vs_oracle(engine, shape):
    rng = np.random.default_rng(sum(shape))
    x = _relu_quant(rng, shape)                       # ties and all-zero patches everywhere
    pooled, idx = engine.maxpool_switch(x)
    p_ref, sw_ref, idx_ref = rs.maxpool_with_switch(x)
    assert np.array_equal(idx, idx_ref)
    assert np.array_equal(pooled, p_ref.astype(np.float32))
    if shape[1] % 2 == 0 and shape[2] % 2 == 0:
        g = rng.normal(0, 1, pooled.shape).astype(np.float32)
        un = engine.unpool_switch(g, idx)
        assert np.array_equal(un.astype(np.float64), rs.unpool_with_switch(g, sw_ref))
        # scatter positions: exactly one non-zero candidate per patch, at the switch
        assert np.count_nonzero(un) <= g.size


def test_pool_empty_and_odd(engine):
    pooled, idx = engine.maxpool_switch(np.zeros((1, 1, 1, 4), np.float32))
    assert pooled.shape == (1, 0, 0, 4)
    from deconv_api_b200 import _native
    import ctypes
    out = np.zeros((1, 5, 4, 2), np.float32)
    x = np.zeros((1, 2, 2, 2), np.float32)
    sw = np.zeros((1, 2, 2, 2), np.uint8)
    rc = engine.lib.dcv_unpool_switch(engine.ctx, x.ctypes.data, sw.ctypes.data, 1, 5, 4, 2, out.ctypes.data)
    assert rc == -1 and b"even" in engine.lib.dcv_last_error(engine.ctx)


@pytest.fixture(params=["tensor_cores", "ffma"])
def fp32_kernels(request, engine):
    """Both kernel families behind the fp32 operators: the split-operand tcgen05 kernels (conv_exact_kernel, the
    default of the forward tape and of the DCV_FP32 chain; GEMM + col2im form for <= 3 output channels) and the
    CUDA-core FFMA kernel.  Both are held to the same 1e-5."""
    tc = request.param == "tensor_cores"
    engine.set_forward_mode(tc)
    engine.set_fp32_mode(tc)
    yield request.param
    engine.set_forward_mode(True)
    engine.set_fp32_mode(True)


@pytest.mark.parametrize("B,H,W,cin,cout,bias,relu", [
    (1, 16, 16, 3, 64, True, True),        # first layer: K = 27 (FFMA in both modes, like the tape)
    (2, 20, 12, 64, 64, True, True),       # N tile 64
    (1, 14, 14, 512, 512, False, True),    # block5 shape: few rows -> N tile 64, 8 k-blocks
    (1, 9, 11, 8, 5, True, False),         # odd sizes, ragged channels, linear (cin < 16: FFMA)
    (1, 33, 17, 16, 24, False, True),      # smallest tensor-core K (padded to 64), ragged N
    (1, 224, 224, 3, 64, True, True),
    (4, 56, 56, 64, 256, True, True),      # enough rows for the N tile 128 form (204 tiles)
    (1, 33, 17, 16, 8, True, True),        # <= 16 output channels: N tile 16
    (2, 9, 7, 40, 72, True, False),        # ragged K and N with bias, odd sizes, linear
    (300, 4, 4, 64, 64, True, True),       # many tiny items: tiles straddle many item boundaries
    (3, 28, 28, 256, 512, False, True),    # block4 shape, batch 3
])
def test_conv3x3_vs_oracle(engine, fp32_kernels, B, H, W, cin, cout, bias, relu):
    rng = np.random.default_rng(B * 1000 + H + cin + cout)
    x = rng.normal(0, 1, (B, H, W, cin)).astype(np.float32)
    lim = np.sqrt(6.0 / (9 * cin + 9 * cout))
    Wt = rng.uniform(-lim, lim, (3, 3, cin, cout)).astype(np.float32)
    b = rng.uniform(-0.1, 0.1, (cout,)).astype(np.float32) if bias else np.zeros((cout,), np.float32)
    out = engine.conv3x3(x, Wt, b, relu=relu)
    ref = rs.conv_up(x, Wt, b, relu=relu)
    assert out.shape == ref.shape
    assert rel_l2(out, ref) < FP32_TOL


@pytest.mark.parametrize("B,H,W,cin,cout,relu", [
    (1, 16, 16, 3, 64, True),              # last backward layer: 64 -> 3 channels (exact GEMM + col2im)
    (2, 12, 20, 64, 128, True),
    (1, 14, 14, 512, 512, True),
    (1, 7, 9, 5, 8, False),
    (2, 31, 45, 3, 64, True),              # last-layer kernel: ragged 14x14 patches, two items
    (4, 56, 56, 128, 64, True),            # N tile 64 with K = 64... the block2_conv1 shape
    (6, 56, 56, 256, 128, True),           # N tile 128 (228 tiles)
    (2, 9, 7, 72, 40, False),              # ragged channels, odd sizes, linear
    (300, 4, 4, 64, 64, True),
    (1, 33, 17, 8, 16, True),              # N tile 16 through the generic kernel (9 * 8 > 32)
    (8, 56, 56, 64, 128, True),            # 64 output channels, K = 128, 204 tiles
    (2, 120, 97, 64, 64, True),            # K = 64, odd sizes (tiles straddle image rows and the two items)
    (40, 47, 45, 48, 64, False),           # ragged output channels (48 of 64), many items, linear
])
def test_deconv3x3_fp32_vs_oracle(engine, fp32_kernels, B, H, W, cin, cout, relu):
    rng = np.random.default_rng(B * 1000 + H + cin + cout + 1)
    x = np.maximum(rng.normal(0, 1, (B, H, W, cout)), 0).astype(np.float32)
    lim = np.sqrt(6.0 / (9 * cin + 9 * cout))
    Wt = rng.uniform(-lim, lim, (3, 3, cin, cout)).astype(np.float32)
    out = engine.deconv3x3(x, Wt, relu=relu, precision="fp32")
    ref = rs.conv_down(x, Wt, relu=relu)
    assert out.shape == (B, H, W, cin)
    assert rel_l2(out, ref) < FP32_TOL
    if relu:
        assert out.min() >= 0


def test_exact_kernels_against_float64(engine):
    """The split-operand tensor-core conv against a float64 convolution of the same fp32 operands: the error of the
    kernel itself (six of nine cross products, fp32 register accumulation), without the oracle's own fp32 rounding."""
    import torch
    import torch.nn.functional as F
    rng = np.random.default_rng(123)
    x = rng.normal(0, 1, (2, 28, 28, 256)).astype(np.float32)
    lim = np.sqrt(6.0 / (9 * 256 + 9 * 256))
    Wt = rng.uniform(-lim, lim, (3, 3, 256, 256)).astype(np.float32)
    ref64 = F.conv2d(torch.from_numpy(x).double().permute(0, 3, 1, 2), torch.from_numpy(Wt).double().permute(3, 2, 0, 1),
                     padding=1).permute(0, 2, 3, 1).numpy()
    engine.set_forward_mode(True)
    out = engine.conv3x3(x, Wt, None, relu=False)
    assert rel_l2(out, ref64) < 2e-6
    engine.set_forward_mode(False)
    ffma = engine.conv3x3(x, Wt, None, relu=False)
    engine.set_forward_mode(True)
    assert rel_l2(ffma, ref64) < 2e-6


def test_channel_sums_and_top_filters(engine):
    from deconv_api_b200 import deepdream as dd
    rng = np.random.default_rng(9)
    out = rng.normal(0.05, 1, (1, 28, 28, 512)).astype(np.float32)
    sums = engine.channel_sums_array(out.reshape(-1, 512))
    ref = out.astype(np.float64).sum(axis=(0, 1, 2))
    assert np.allclose(sums, ref, rtol=1e-12, atol=1e-9)
    top = dd.find_top_filters(out)
    top_ref = rs.find_top_filters(out)
    assert [t[0] for t in top] == [t[0] for t in top_ref]
    assert np.allclose([t[1] for t in top], [t[1] for t in top_ref], rtol=1e-5)
    assert isinstance(top[0][0], int) and len(top) == 8
    # 2-D (dense) outputs and the s > 0 rule
    o2 = np.array([[1.0, -2.0, 0.0, 3.0]], np.float32)
    assert [t[0] for t in dd.find_top_filters(o2)] == [3, 0]


def test_deprocess_golden(engine, golden):
    from deconv_api_b200 import deepdream as dd
    g = golden("deprocess_reference.npz")
    x = g["small_in"].copy()
    out = dd.deprocess_image(x)
    assert out.dtype == np.uint8 and out.shape == x.shape
    d = np.abs(out.astype(np.int16) - g["small_u8"].astype(np.int16))
    assert d.max() <= 1 and np.count_nonzero(d) <= out.size // 1000      # truncation boundary cases only
    assert np.allclose(x, g["small_mutated"], rtol=0, atol=2e-6)         # in-place side effect reproduced
    # large, odd length, constant input (std = 0 -> epsilon path)
    big = np.random.default_rng(1).normal(2.0, 3.0, (448 * 448 * 3 + 3,)).astype(np.float32)
    ref = rs.deprocess_image(big.copy())
    got = engine.deprocess(big.copy(), mutate=False)
    d = np.abs(got.astype(np.int16) - ref.astype(np.int16))
    assert d.max() <= 1 and np.count_nonzero(d) <= got.size // 1000
    const = np.full((64,), 3.0, np.float32)
    assert np.array_equal(engine.deprocess(const.copy()), rs.deprocess_image(const.copy()))


def test_dense_and_activation_ops(engine):
    rng = np.random.default_rng(4)
    x = rng.normal(0, 1, (2, 300)).astype(np.float32)
    W = rng.normal(0, 0.1, (300, 70)).astype(np.float32)
    b = rng.normal(0, 0.1, (70,)).astype(np.float32)
    y = engine.dense(x, W, b, "relu")
    assert rel_l2(y, np.maximum(x @ W + b, 0)) < FP32_TOL
    s = engine.dense(x, W, b, "softmax")
    assert rel_l2(s, rs.softmax(x @ W + b)) < FP32_TOL
    g = rng.normal(0, 1, (2, 70)).astype(np.float32)
    assert rel_l2(engine.dense_transpose(g, W), g @ W.T) < FP32_TOL
    assert np.array_equal(engine.relu(x), np.maximum(x, 0))


def test_wrapper_classes_mirror_reference(engine, golden):
    """The D* classes keep the reference's contract: ctor takes a layer, .up/.down, up_data/down_data/switch."""
    from deconv_api_b200 import deepdream as dd, layers
    g = golden("pool_reference.npz")
    p = dd.DPooling(layers.MaxPooling2D((2, 2), name="p"))
    up = p.up(g["x"])
    assert up.dtype == np.float64 and np.array_equal(up, g["pooled"])
    assert p.switch.dtype == np.float64 and np.array_equal(p.switch.astype(np.uint8), g["switch"])
    down = p.down(g["g"])
    assert np.array_equal(down, g["unpooled"]) and np.array_equal(p.down_data, down)
    conv = layers.Convolution2D(8, (3, 3), activation="relu", padding="same", name="c")
    conv._build((None, 10, 10, 4))
    rng = np.random.default_rng(0)
    Wt = rng.normal(0, 0.2, (3, 3, 4, 8)).astype(np.float32)
    b = rng.normal(0, 0.1, (8,)).astype(np.float32)
    conv.set_weights([Wt, b])
    c = dd.DConvolution2D(conv)
    x = rng.normal(0, 1, (1, 10, 10, 4)).astype(np.float32)
    assert rel_l2(c.up(x), rs.conv_up(x, Wt, b)) < FP32_TOL
    assert rel_l2(c.down(c.up_data), rs.conv_down(c.up_data, Wt)) < FP32_TOL
    a = dd.DActivation(conv)
    assert np.array_equal(a.up(x), np.maximum(x, 0)) and np.array_equal(a.down(-x), np.maximum(-x, 0))
    i = dd.DInput(layers.InputLayer((10, 10, 4)))
    assert i.up(x) is x and i.down(x) is x


def test_error_paths(engine):
    """Call-order and argument errors come back as negative status codes with a message, never as a crash."""
    import ctypes
    from deconv_api_b200 import engine as E, _native
    e = E.Engine(0)
    try:
        lib = e.lib
        feats = np.zeros(1, np.int32)
        out = np.zeros((1, 1, 8, 8, 3), np.float32)
        # project before any network / forward
        assert lib.dcv_project(e.ctx, 0, feats.ctypes.data, 1, 0, 0, out.ctypes.data) < 0
        W = np.zeros((3, 3, 3, 4), np.float32)
        e.add_conv("c", W, None)
        assert lib.dcv_project(e.ctx, 0, feats.ctypes.data, 1, 0, 0, out.ctypes.data) == -3      # DCV_ERR_STATE: no tape
        assert b"dcv_forward" in lib.dcv_last_error(e.ctx)
        e.forward(np.zeros((1, 8, 8, 3), np.float32), 0)
        bad = np.array([7], np.int32)
        assert lib.dcv_project(e.ctx, 0, bad.ctypes.data, 1, 0, 0, out.ctypes.data) == -1         # feature out of range
        assert lib.dcv_project(e.ctx, 0, feats.ctypes.data, 1, 5, 0, out.ctypes.data) == -1       # illegal visualize mode
        assert lib.dcv_project(e.ctx, 0, feats.ctypes.data, 1, 0, 9, out.ctypes.data) == -1       # unknown precision
        assert lib.dcv_forward(e.ctx, None, 1, 8, 8, 0) == -1
        assert lib.dcv_forward(e.ctx, out.ctypes.data, 1, 8, 8, 3) == -1                          # layer out of range
        with pytest.raises(_native.DcvError):
            e.forward(np.zeros((1, 8, 8, 5), np.float32), 0)                                      # wrong channel count
        assert lib.dcv_set_forward_mode(e.ctx, 7) == -1
        assert e.project(0, [0, 1]).shape == (2, 1, 8, 8, 3)
    finally:
        e.close()


@pytest.mark.gpu
def test_device_resident_ops_match_host_forms(engine):
    """dcv_*_dev entry points (caller-owned device buffers) give the same bits as the host-buffer forms, which
    are the ones checked against the oracle above; resize also at shapes that exercise every branch of the strip
    kernel (enlarging, shrinking, identity, ragged last strip / last block)."""
    import torch
    from oracle import dream_oracle as do
    rng = np.random.default_rng(5)
    dev = torch.device("cuda:0")
    x = np.maximum(rng.normal(0, 1, (3, 10, 12, 24)), 0).astype(np.float32)
    x[0, :2, :2, :] = 0.0                                   # all-equal patches: switch (0,0)
    pooled, sw = engine.maxpool_switch(x)
    xd = torch.from_numpy(x).to(dev)
    pd = torch.empty((3, 5, 6, 24), device=dev)
    sd = torch.empty((3, 5, 6, 24), device=dev, dtype=torch.uint8)
    torch.cuda.synchronize()
    engine.maxpool_switch_dev(xd.data_ptr(), x.shape, pd.data_ptr(), sd.data_ptr())
    engine.sync()
    assert np.array_equal(pd.cpu().numpy(), pooled) and np.array_equal(sd.cpu().numpy(), sw)
    ud = torch.empty_like(xd)
    engine.unpool_switch_dev(pd.data_ptr(), sd.data_ptr(), x.shape, ud.data_ptr())
    engine.sync()
    assert np.array_equal(ud.cpu().numpy(), engine.unpool_switch(pooled, sw))
    with pytest.raises(Exception):
        engine.unpool_switch_dev(pd.data_ptr(), sd.data_ptr(), (3, 9, 12, 24), ud.data_ptr())     # odd output size

    for (h, w, c), (oh, ow) in (((37, 53, 3), (61, 44)), ((64, 64, 3), (45, 45)), ((100, 90, 3), (140, 126)),
                                ((20, 30, 5), (20, 30)), ((9, 7, 1), (33, 300)), ((300, 260, 3), (17, 19))):
        img = rng.normal(0, 50, (2, h, w, c)).astype(np.float32)
        idv = torch.from_numpy(img).to(dev)
        od = torch.empty((2, oh, ow, c), device=dev)
        torch.cuda.synchronize()
        engine.resize_bilinear_dev(idv.data_ptr(), img.shape, (oh, ow), od.data_ptr())
        engine.sync()
        got = od.cpu().numpy()
        assert np.array_equal(got, engine.resize_bilinear(img, (oh, ow)))
        assert rel_l2(got, do.resize(img, (oh, ow))) < 1e-6
    img = rng.normal(0, 1, (2, 33, 700, 3)).astype(np.float32)       # 2100 floats per row: unrolled body + tail
    idv = torch.from_numpy(img).to(dev)
    od = torch.empty_like(idv)
    torch.cuda.synchronize()
    for oy, ox in ((0, 0), (5, -7), (-32, 32), (33, 700), (40, -1401)):
        engine.roll_dev(idv.data_ptr(), img.shape, oy, ox, od.data_ptr())
        engine.sync()
        assert np.array_equal(od.cpu().numpy(), np.roll(img, (oy, ox), axis=(1, 2)))
    with pytest.raises(Exception):
        engine.roll_dev(idv.data_ptr(), img.shape, 1, 1, idv.data_ptr())                          # in place


_DEPROCESS_SCRIPT = r"""
import sys, numpy as np
from deconv_api_b200 import engine as E
eng = E.Engine(0)
rng = np.random.default_rng(11)
out = {}
n = 3 * 2048 * 37 + 3                       # ragged: a partial last chunk and n % 4 != 0
cases = {
    "normal": rng.normal(2.0, 3.0, n),
    "wide": rng.normal(0, 1, n) * 10.0 ** rng.uniform(-30, 30, n),          # 60 decades of magnitude
    "tiny_std": 1.0 + rng.normal(0, 1e-6, n),
    "huge": rng.normal(0, 1e37, n),
    "subnormal": rng.normal(0, 1e-41, n),
    "ones_mantissa": np.ldexp(2.0 - 2.0 ** -23, rng.integers(-20, 20, n)),  # divisor-like worst cases as dividends
}
for k, v in cases.items():
    x = v.astype(np.float32)
    out[k + "_u8"] = eng.deprocess(x, mutate=True)
    out[k + "_norm"] = x
np.savez(sys.argv[1], **out)
"""


def test_deprocess_pipelined_map_equals_division_form(tmp_path):
    """The pipelined map pass (bulk async copies, quotient by reciprocal + two FMA corrections) against the grid-stride
    form that divides (DCV_DEPROCESS_PIPE=0): uint8 output AND the normalised floats left in the argument are equal
    bit for bit, on ordinary data and on data built to stress the quotient (60 decades of magnitude, std -> 0,
    values next to overflow, subnormals, all-ones mantissas)."""
    import os, subprocess, sys
    res = {}
    for mode in ("0", "1"):
        path = str(tmp_path / ("d%s.npz" % mode))
        env = dict(os.environ, DCV_DEPROCESS_PIPE=mode, PYTHONPATH=os.pathsep.join(sys.path))
        subprocess.run([sys.executable, "-W", "ignore", "-c", _DEPROCESS_SCRIPT, path], check=True, env=env, timeout=300)
        res[mode] = np.load(path)
    for k in res["0"].files:
        a, b = res["0"][k], res["1"][k]
        assert np.array_equal(a.view(np.uint8 if a.dtype == np.uint8 else np.uint32), b.view(np.uint8 if b.dtype == np.uint8 else np.uint32)), k
