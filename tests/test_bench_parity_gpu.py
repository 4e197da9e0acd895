"""Parity of the configurations that bench.py actually times (BASELINE.json configs[1] and the shape of configs[3]),
through the C-ABI, against the oracle — every projection, both precisions.

  * configs[1]: one 224x224 image, forward to block4_conv3, ALL 512 feature maps projected;
  * configs[3] shape: a batch of VGG-size images sharing one tape, block5_conv3 on top, 64 maps: every (feature, image)
    item reads the switches of ITS image (item % B in the kernels).

Tolerances (north_star): rel-L2 <= 1e-5 per projection in the fp32 mode, <= 2e-2 in the bf16 mode.  Two levels:
  * the backward chain of EVERY map is held to the tolerance given the tape (the oracle follows the library's
    switches where they are near-ties, and is seeded with the library's top activations after those were checked
    against its own: conftest.following_tape);
  * end to end, oracle's own forward pass included: all 512 projections taken together within the tolerance, and per
    map the distribution profiles/diag_512.py prints (fp32: median 3e-6, 99th percentile 1.8e-5, worst 2e-4 on a map
    with ONE non-zero activation of 6.8e-3).  A projection is linear in its seed, the seed is the forward activation,
    and two fp32 forward passes agree to ~1e-6 of the LAYER's scale: a weak map carries that as a larger relative
    error whichever implementation is "right".  Hence per map: 90 % of the live maps within 1e-5, every map whose
    peak activation is >= 10 % of the layer's within 3e-5 (fp32); the bf16 mode keeps 2e-2 for every live map."""
import numpy as np
import pytest

from oracle import keras_shim
from deconv_api_b200 import synth, vgg16
from conftest import following_tape, rel_l2

pytestmark = pytest.mark.gpu

FP32_TOL, BF16_TOL = 1e-5, 2e-2


def _project_ref(tape, feats, batch=32):
    out = []
    for i in range(0, len(feats), batch):
        out.append(tape.project([int(f) for f in feats[i:i + batch]]))
    return np.concatenate(out, axis=0)


def test_all_512_block4_conv3_maps_vs_oracle():
    """The benchmarked step itself: bench.py's image (seed 0), every one of the 512 maps, fp32 and bf16."""
    model = vgg16.VGG16(weights=None, include_top=False, input_shape=(224, 224, 3))
    ref_model = keras_shim.VGG16(include_top=False, input_shape=(224, 224, 3))
    img = synth.image_batch(1, 224, 224, seed0=0)
    eng = model.bound_engine(upto="block4_conv3")
    li = eng.index["block4_conv3"]
    eng.forward(img, li)
    feats = np.arange(512, dtype=np.int32)
    own, _ = following_tape(eng, ref_model, "block4_conv3", img)                       # oracle's own forward values
    want_own = _project_ref(own, feats)
    tape, _ = following_tape(eng, ref_model, "block4_conv3", img, follow_top=True)     # seeded with the library's tape
    want = _project_ref(tape, feats)
    act = eng.activation(li)[0]
    peak = act.max(axis=(0, 1))
    live = [k for k in range(512) if peak[k] > 0]
    strong = [k for k in live if peak[k] >= 0.1 * peak.max()]
    assert len(live) > 400 and len(strong) > 300              # random-init VGG16: most maps are alive
    for prec, tol in (("fp32", FP32_TOL), ("bf16", BF16_TOL)):
        got = eng.project(li, feats, precision=prec)[:, 0]
        assert got.shape == want.shape == (512, 224, 224, 3)
        worst = max(rel_l2(got[k], want[k]) for k in live)
        assert worst < tol, (prec, "chain", worst)
        e2e = {k: rel_l2(got[k], want_own[k]) for k in live}
        assert rel_l2(got, want_own) < tol, (prec, "all 512 together")
        if prec == "fp32":
            assert np.percentile(list(e2e.values()), 90) < tol, (prec, "end to end, 90th percentile")
            assert max(e2e[k] for k in strong) < 3 * tol, (prec, "end to end, strong maps", max(e2e[k] for k in strong))
        else:
            assert max(e2e.values()) < tol, (prec, "end to end", max(e2e.values()))
        dead = [k for k in range(512) if k not in live]
        assert all(np.abs(got[k]).max() == 0 for k in dead)   # a dead map projects to exactly zero
        assert got.min() >= 0                                 # every projection ends in block1_conv1's own ReLU


def test_config4_shape_batch_of_images_block5_conv3():
    """Two VGG-size images in ONE tape, block5_conv3 on top, 64 maps, both precisions, per image against the oracle.
    Image seed 4 has a 1-ulp near-tie in block2_pool (profiles/diag_b2.py): the flip handling is exercised too."""
    model = vgg16.VGG16(weights=None, include_top=False, input_shape=(224, 224, 3))
    ref_model = keras_shim.VGG16(include_top=False, input_shape=(224, 224, 3))
    img = synth.image_batch(2, 224, 224, seed0=3)
    eng = model.bound_engine(upto="block5_conv3")
    li = eng.index["block5_conv3"]
    eng.forward(img, li)
    sums = eng.channel_sums(li)
    feats = np.argsort(-sums, kind="stable")[:48].astype(np.int32)
    feats = np.concatenate([feats, np.arange(0, 512, 32, dtype=np.int32)])          # + 16 fixed ones (some weak / dead)
    assert feats.size == 64
    tapes = [following_tape(eng, ref_model, "block5_conv3", img[b:b + 1], b=b, follow_top=True)[0] for b in range(2)]
    want = np.stack([_project_ref(t, feats) for t in tapes], axis=1)              # (64, 2, 224, 224, 3)
    for prec, tol in (("fp32", FP32_TOL), ("bf16", BF16_TOL)):
        got = eng.project(li, feats, precision=prec)
        assert got.shape == want.shape
        for k in range(64):
            for b in range(2):
                if np.abs(want[k, b]).max() > 0:
                    assert rel_l2(got[k, b], want[k, b]) < tol, (prec, k, b)
                else:
                    assert np.abs(got[k, b]).max() == 0
    # the two images really are different units: image 1's projections are not image 0's
    assert rel_l2(want[0, 1], want[0, 0]) > 0.1


def test_batched_tape_equals_single_image_tapes_at_full_size():
    """Size-independent property for the configs[3] path at its real size: 8 images in one tape (what
    parallel.visualize_batch runs per block) must give, image by image, what 8 single-image tapes give — every item
    reads its own image's switches and activations.  bf16 chain, uint8 and float32 results."""
    model = vgg16.VGG16(weights=None, include_top=False, input_shape=(224, 224, 3))
    imgs = synth.image_batch(8, 224, 224, seed0=10)
    eng = model.bound_engine(upto="block5_conv3")
    li = eng.index["block5_conv3"]
    feats = np.arange(0, 512, 16, dtype=np.int32)                  # 32 maps x 8 images = 256 items
    eng.forward(imgs, li)
    both = eng.project(li, feats, precision="bf16")
    both8 = eng.project(li, feats, precision="bf16", out_format="u8")
    assert both.shape == (32, 8, 224, 224, 3)
    for b in (0, 3, 7):
        eng.forward(imgs[b:b + 1], li)
        one = eng.project(li, feats, precision="bf16")[:, 0]
        one8 = eng.project(li, feats, precision="bf16", out_format="u8")[:, 0]
        # the batch decides tile shapes (more rows: other kernel variants), so bits may differ at bf16 level
        assert rel_l2(both[:, b], one) < 5e-3, b
        d = np.abs(both8[:, b].astype(np.int16) - one8.astype(np.int16))
        assert np.percentile(d, 99.9) <= 2, (b, d.max())
