"""Unit parity of the fused 64-channel tail kernel (csrc/tc_strip.cuh) through the C-ABI (dcv_deconv_tail), all four
forms: plain strip conv, unpool fused in front (DPooling.down, deepdream.py:191-209), last conv fused behind
(DConvolution2D.down of block1_conv1, deepdream.py:102-111), both.

Tolerances.  Against the oracle evaluated on the SAME bf16-rounded operands (and with the 64-channel intermediate
rounded to bf16 where the kernel rounds it) only the fp32 summation order differs, plus a rare one-ulp flip where a
sum sits on a bf16 rounding boundary: rel-L2 <= 1e-3 (measured ~1e-5 .. 1e-4) — this pins indexing, swizzle, strip
halos, switch handling and the col2im exactly.  Against the unrounded fp32 oracle north_star's bf16 bound: 2e-2."""
import numpy as np
import pytest
import torch

from oracle import restatement as rs
from conftest import rel_l2

pytestmark = pytest.mark.gpu


def bf16_round(a):
    return torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(torch.bfloat16).to(torch.float32).numpy()


def glorot(rng, cin, cout):
    lim = np.sqrt(6.0 / (9 * cin + 9 * cout))
    return rng.uniform(-lim, lim, (3, 3, cin, cout)).astype(np.float32)


@pytest.mark.parametrize("unpool", [False, True])
@pytest.mark.parametrize("last", [False, True])
@pytest.mark.parametrize("B,H,W", [
    (2, 32, 32),            # one column block, strips of 8 tall rows, two items
    (1, 224, 224),          # VGG16 block1: two column blocks
    (3, 112, 112),          # the 112 x 112 layers: one block of 113 positions
    (1, 20, 120),           # w + 1 = 121: second block starts at 118
    (5, 6, 250),            # three column blocks, short images (item boundaries inside a strip)
])
def test_tail_vs_oracle(engine, unpool, last, B, H, W):
    rng = np.random.default_rng(1000 * B + H + W + 2 * unpool + last)
    ih, iw = (H // 2, W // 2) if unpool else (H, W)
    x = np.maximum(rng.normal(0, 1, (B, ih, iw, 64)), 0).astype(np.float32)
    x[rng.random(x.shape) < 0.3] = 0
    sw = rng.integers(0, 4, x.shape).astype(np.uint8) if unpool else None
    W2 = glorot(rng, 64, 64)
    W1 = glorot(rng, 3, 64) if last else None
    got = engine.deconv_tail(x, W2, W1, sw)
    assert got.shape == (B, H, W, 3 if last else 64)

    def chain(xx, w2, w1, rnd):
        u = rs.unpool_with_idx(xx, sw) if unpool else xx
        m = rs.conv_down(u, w2, relu=True)
        if w1 is None:
            return m
        return rs.conv_down(rnd(m), w1, relu=True)
    exact = chain(bf16_round(x), bf16_round(W2), bf16_round(W1) if last else None, bf16_round)
    if not last:
        exact = bf16_round(exact)                     # the kernel's output is bf16
    assert rel_l2(got, exact) < 1e-3, rel_l2(got, exact)
    full = chain(x, W2, W1, lambda a: a)
    assert rel_l2(got, full) < 2e-2
    assert got.min() >= 0


def test_tail_linear_and_unsupported(engine):
    rng = np.random.default_rng(5)
    x = rng.normal(0, 1, (1, 16, 16, 64)).astype(np.float32)
    W2 = glorot(rng, 64, 64)
    got = engine.deconv_tail(x, W2, None, None, relu2=False)
    exact = bf16_round(rs.conv_down(bf16_round(x), bf16_round(W2), relu=False))
    assert rel_l2(got, exact) < 1e-3 and got.min() < 0
    from deconv_api_b200 import _native
    with pytest.raises(_native.DcvError):
        engine.deconv_tail(rng.normal(0, 1, (1, 8, 8, 128)).astype(np.float32), glorot(rng, 64, 128))


def test_fused_tail_equals_the_three_operators_it_replaces_at_full_size(engine):
    """Size-independent property at the benchmark's own size (224 x 224, 64 channels): the fused kernel against the
    chain of stand-alone operators it replaces — DPooling.down, then DConvolution2D.down of block1_conv2 and of
    block1_conv1 in the bf16 mode (each of those is pinned against the oracle on its own in test_tc_gpu.py /
    test_ops_gpu.py).  Same bf16 roundings at the same places, so only the fp32 summation order differs."""
    rng = np.random.default_rng(77)
    B = 4
    x = np.maximum(rng.normal(0, 1, (B, 112, 112, 64)), 0).astype(np.float32)
    x[rng.random(x.shape) < 0.5] = 0
    sw = rng.integers(0, 4, x.shape).astype(np.uint8)
    W2, W1 = glorot(rng, 64, 64), glorot(rng, 3, 64)
    fused = engine.deconv_tail(x, W2, W1, sw)
    step1 = engine.unpool_switch(bf16_round(x), sw)
    step2 = engine.deconv3x3(step1, W2, relu=True, precision="bf16")
    step3 = engine.deconv3x3(bf16_round(step2), W1, relu=True, precision="bf16")
    assert fused.shape == step3.shape == (B, 224, 224, 3)
    assert rel_l2(fused, step3) < 1e-3, rel_l2(fused, step3)
    # and the half-fused forms agree with it too
    only_unpool = engine.deconv_tail(x, W2, None, sw)
    assert rel_l2(only_unpool, bf16_round(step2)) < 1e-3
    only_last = engine.deconv_tail(step1, W2, W1, None)
    assert rel_l2(only_last, step3) < 1e-3
