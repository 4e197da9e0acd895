"""Compact result formats of the projection entry point (dcv_project_fmt) and the fused request tail
(dcv_request_mosaic), through the C-ABI, against the oracle.

uint8 results are `deprocess_image` (reference deepdream.py:483-498) applied to each projection — what the
reference's only consumer does with them (main.py:72).  Stated tolerance for bytes: at most one level of difference
on at most 0.2 % of the pixels (the truncation at astype('uint8') amplifies the last-ulp difference between the
GPU's fp64 mean/std and numpy's pairwise-fp32 ones); bf16 results must be the round-to-nearest-even of the fp32
results bit for bit."""
import numpy as np
import pytest
import torch

from oracle import keras_shim, restatement as rs
from deconv_api_b200 import deepdream as dd, engine as E, synth, vgg16
from conftest import following_tape, rel_l2

pytestmark = pytest.mark.gpu


def bytes_close(got, ref, frac=0.002):
    d = np.abs(got.astype(np.int16) - ref.astype(np.int16))
    return d.max() <= 1 and np.count_nonzero(d) <= max(1, int(frac * d.size))


def oracle_u8(proj_f32):
    """deprocess_image of the oracle on every (feature, image) projection of an (n,B,H,W,C) array."""
    out = np.empty(proj_f32.shape, np.uint8)
    for k in range(proj_f32.shape[0]):
        for b in range(proj_f32.shape[1]):
            out[k, b] = rs.deprocess_image(proj_f32[k, b].copy())
    return out


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_project_u8_and_bf16_formats_vgg16(precision):
    model = vgg16.VGG16(weights=None, include_top=False, input_shape=(224, 224, 3))
    ref_model = keras_shim.VGG16(include_top=False, input_shape=(224, 224, 3))
    img = synth.image_batch(2, 224, 224, 3)                      # B = 2: statistics are per (feature, image)
    eng = model.bound_engine(upto="block3_conv2")
    li = eng.index["block3_conv2"]
    eng.forward(img, li)
    feats = [5, 17, 100, 255, 3, 42]
    f32 = eng.project(li, feats, precision=precision)
    u8 = eng.project(li, feats, precision=precision, out_format="u8")
    b16 = eng.project(li, feats, precision=precision, out_format="bf16")
    assert u8.shape == f32.shape and u8.dtype == np.uint8 and b16.dtype == np.uint16
    # bf16: exactly the RNE rounding of the fp32 result
    want = torch.from_numpy(f32).to(torch.bfloat16).view(torch.uint16).numpy()
    assert np.array_equal(b16, want)
    assert np.array_equal(E.bf16_bits_to_f32(b16), torch.from_numpy(f32).to(torch.bfloat16).float().numpy())
    # u8 against deprocess_image of the library's own fp32 projections ...
    assert bytes_close(u8, oracle_u8(f32))
    # ... and end to end against the oracle's projections (fp32 mode: same pixels up to the 1e-5 chain tolerance)
    if precision == "fp32":
        # image seed 4 has a 1-ulp near-tie in block2_pool (profiles/diag_b2.py)
        tapes = [following_tape(eng, ref_model, "block3_conv2", img[b:b + 1], b=b)[0] for b in range(2)]
        ref = np.stack([np.stack([t.project([f])[0] for t in tapes]) for f in feats])
        assert rel_l2(f32, ref) < 1e-5
        assert bytes_close(u8, oracle_u8(ref), frac=0.004)
    # every projection got its own statistics: a constant-free sanity check on the byte histogram
    assert all(40 < int(u8[k, b].mean()) < 215 for k in range(len(feats)) for b in range(2))


def test_project_u8_pipelined_host_output_and_ragged_item_length():
    """n >= 64 takes the chunked, D2H-pipelined path (tapered chunks, two staging buffers); a 9x7 image gives
    items of 189 floats (not a multiple of 4: the scalar kernels)."""
    model = vgg16.VGG16(weights=None, include_top=False, input_shape=(None, None, 3), seed=3, bias=True)
    eng = model.bound_engine(upto="block1_conv2")
    li = eng.index["block1_conv2"]
    for hw in ((32, 32), (9, 7)):
        img = synth.caffe_preprocess(synth.smooth_image_u8(2, *hw))
        eng.forward(img, li)
        feats = list(range(64))
        f32 = eng.project(li, feats)
        u8 = eng.project(li, feats, out_format="u8")
        assert bytes_close(u8, oracle_u8(f32), frac=0.01)
        dev = torch.empty(f32.size, dtype=torch.uint8, device="cuda")
        eng.project_dev(li, feats, dev.data_ptr(), out_format="u8")
        eng.sync()
        assert np.array_equal(dev.cpu().numpy().reshape(u8.shape), u8)      # device and host outputs agree exactly
        b16 = eng.project(li, feats, out_format="bf16")
        assert np.array_equal(b16, torch.from_numpy(f32).to(torch.bfloat16).view(torch.uint16).numpy())


@pytest.mark.parametrize("layer", ["block3_conv3", "block2_pool", "block1_conv2"])
def test_request_mosaic_equals_the_step_by_step_tail(layer):
    """dcv_request_mosaic against the same request done the reference's way, step by step (main.py:60-72):
    preprocess_input, forward, find_top_filters, the four projections, np.concatenate, deprocess_image — (a) with the
    library's own operators for each step (fused vs unfused: same bytes up to the de-process tolerance), (b) against the
    oracle (preprocess exact, top-4 identical, projections 1e-5 with near-tie switches followed)."""
    model = vgg16.VGG16(weights=None, include_top=True)
    ref_model = keras_shim.VGG16(include_top=True)
    img_u8 = synth.smooth_image_u8(5, 224, 224)[..., ::-1].copy()         # BGR uint8 as cv2 hands it over
    eng = model.bound_engine(upto=layer)
    li = eng.index[layer]
    mosaic, feats = eng.request_mosaic(img_u8, li)
    assert mosaic.shape == (448, 448, 3) and mosaic.dtype == np.uint8
    x = vgg16.preprocess_input(np.expand_dims(img_u8.astype(np.float32), 0))
    assert np.array_equal(x, keras_shim.preprocess_input(np.expand_dims(img_u8.astype(np.float32), 0)))
    # (a) unfused, same library
    eng.forward(x, li)
    top = dd._top_from_sums(eng.channel_sums(li), 8, np.float64 if layer.endswith("pool") else np.float32)
    assert [f for f, _ in top[:4]] == list(feats)
    proj = eng.project(li, feats)[:, 0]
    concat = np.concatenate((np.concatenate((proj[0], proj[1]), axis=1), np.concatenate((proj[2], proj[3]), axis=1)), axis=0)
    assert bytes_close(mosaic, rs.deprocess_image(concat.copy()))
    # (b) the oracle: same four maps, same projections
    entries = rs.forward(rs.build_entries(ref_model, layer), x)
    ref_top = rs.find_top_filters(entries[[e.name for e in entries].index(layer)].up)
    assert [f for f, _ in ref_top[:4]] == list(feats)
    tape, _ = following_tape(eng, keras_shim.VGG16(include_top=False, input_shape=(224, 224, 3)), layer, x)
    want = tape.project([int(f) for f in feats])
    assert rel_l2(proj, want) < 1e-5
    # the tape is left valid: a follow-up projection of the same request needs no second forward
    again = eng.project(li, feats[:1])
    assert np.array_equal(again[0, 0], proj[0])


def test_request_mosaic_with_fewer_than_four_positive_maps_raises_index_error():
    """main.py:67-68 indexes [0..3] of find_top_filters' list; a layer with < 4 positive-sum maps is an IndexError."""
    model = vgg16.VGG16(weights=None, include_top=False, input_shape=(32, 32, 3), seed=4)
    l0 = model.get_layer("block1_conv1")
    W, b = l0.get_weights()
    b[:] = -1e6                                                # ReLU kills every map ...
    b[7] = b[9] = b[30] = 1.0                                  # ... but three
    l0.set_weights([W, b])
    eng = model.bound_engine(upto="block1_conv1")
    with pytest.raises(IndexError):
        eng.request_mosaic(synth.smooth_image_u8(1, 32, 32), eng.index["block1_conv1"])


def test_result_format_and_request_tail_error_paths(engine):
    """Argument checking of the new entry points: every refusal is a negative status with a message, nothing crashes."""
    import ctypes
    lib = engine.lib
    model = vgg16.VGG16(weights=None, include_top=False, input_shape=(32, 32, 3), seed=2)
    eng = model.bound_engine(upto="block1_conv2")
    li = eng.index["block1_conv2"]
    eng.forward(synth.caffe_preprocess(synth.smooth_image_u8(1, 32, 32)), li)
    feats = np.array([0, 1], np.int32)
    out = np.zeros((2, 1, 32, 32, 3), np.uint8)
    P = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    assert lib.dcv_project_fmt(eng.ctx, li, P(feats), 2, 0, 0, 7, P(out)) == -1          # unknown format
    assert b"format" in lib.dcv_last_error(eng.ctx)
    assert lib.dcv_project_fmt(eng.ctx, li, None, 2, 0, 0, 2, P(out)) == -1              # no feature list
    assert lib.dcv_project_fmt(eng.ctx, li, P(feats), 2, 0, 0, 2, None) == -1            # no output
    assert lib.dcv_project_fmt(eng.ctx, li, P(np.array([0, 64], np.int32)), 2, 0, 0, 2, P(out)) == -1   # feature out of range
    assert lib.dcv_project_fmt(eng.ctx, li, P(feats), 0, 0, 0, 2, P(out)) == 0           # nothing to do
    assert lib.dcv_project_fmt(eng.ctx, li + 5, P(feats), 2, 0, 0, 2, P(out)) == -1      # layer outside the network
    img = np.zeros((32, 32, 3), np.uint8)
    mosaic = np.zeros((64, 64, 3), np.uint8)
    assert lib.dcv_request_mosaic(eng.ctx, None, 32, 32, li, 0, 0, P(mosaic), None) == -1
    assert lib.dcv_request_mosaic(eng.ctx, P(img), 0, 32, li, 0, 0, P(mosaic), None) == -1
    assert lib.dcv_request_mosaic(eng.ctx, P(img), 32, 32, 99, 0, 0, P(mosaic), None) == -1
    assert lib.dcv_request_mosaic(eng.ctx, P(img), 32, 32, li, 5, 0, P(mosaic), None) == -1   # illegal visualize mode
    # page-locked buffers: round trip through a projection, then release
    buf = eng.pinned_empty((2, 1, 32, 32, 3), np.uint8)
    got = eng.project(li, feats, out=buf, out_format="u8")
    assert np.array_equal(got, eng.project(li, feats, out_format="u8"))
    p = ctypes.c_void_p()
    assert lib.dcv_host_alloc(eng.ctx, 1 << 20, None) == -1
    assert lib.dcv_host_alloc(eng.ctx, 1 << 20, ctypes.byref(p)) == 0 and p.value
    assert lib.dcv_host_free(eng.ctx, p) == 0
    assert lib.dcv_host_free(eng.ctx, None) == 0
