"""CPU tests of the host-side logic: work partition, weight packing/broadcast over gloo (world size 2),
model description objects, switch mask re-encoding, FastTape (the timed CPU baseline) vs the faithful
restatement."""
import os
import subprocess
import sys

import numpy as np
import pytest

from deconv_api_b200 import deepdream as dd, layers, parallel, synth, vgg16
from oracle import keras_shim, restatement as rs
from conftest import rel_l2

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_partition_covers_everything():
    for n in (0, 1, 7, 64, 512):
        for world in (1, 2, 3, 8):
            got = []
            for r in range(world):
                lo, hi = parallel.partition(n, world, r)
                assert 0 <= lo <= hi <= n and hi - lo in (n // world, n // world + 1)
                got += list(range(lo, hi))
            assert got == list(range(n))


def test_shard_work_by_image_then_feature():
    # config 4: 64 images x 512 maps on 8 GPUs -> 8 images each, all features, no exchange
    units = [parallel.shard_work(64, 512, 8, r) for r in range(8)]
    assert all(len(u) == 8 and all(tuple(f) == (0, 512) for _, *f in u) for u in units)
    assert sorted(i for u in units for i, _, _ in u) == list(range(64))
    # config 2: 1 image on 4 GPUs -> forward replicated, features split contiguously
    units = [parallel.shard_work(1, 512, 4, r) for r in range(4)]
    assert [u[0][1:] for u in units] == [(0, 128), (128, 256), (256, 384), (384, 512)]


def test_model_description_matches_keras_shapes():
    m = vgg16.VGG16(weights=None, include_top=True)
    assert [l.name for l in m.layers] == [n for _, n in synth.VGG16_TOPOLOGY]
    assert m.get_layer("block5_conv3").output_shape == (None, 14, 14, 512)
    assert m.get_layer("flatten").output_shape == (None, 25088)
    assert m.get_layer("block1_pool").pool_size == (2, 2)
    W, b = m.get_layer("block3_conv1").get_weights()
    assert W.shape == (3, 3, 128, 256) and b.shape == (256,)
    ref = keras_shim.VGG16()
    Wr, br = ref.layers[7].get_weights()
    assert ref.layers[7].name == "block3_conv1" and np.array_equal(W, Wr) and np.array_equal(b, br)
    cfg = m.get_layer("block1_conv1").get_config()
    assert cfg["activation"] == "relu" and cfg["padding"] == "same" and cfg["filters"] == 64
    x = np.arange(24, dtype=np.float32).reshape(1, 2, 4, 3)
    y = vgg16.preprocess_input(x.copy())
    assert np.allclose(y[0, 0, 0], [2 - 103.939, 1 - 116.779, 0 - 123.68])


def test_switch_mask_roundtrip(golden):
    g = golden("pool_reference.npz")
    mask = dd.switch_mask(g["idx"], g["x"].shape)
    assert np.array_equal(mask.astype(np.uint8), g["switch"])
    assert np.array_equal(dd.switch_index_from_mask(mask), g["idx"])


def test_weight_pack_roundtrip():
    a = vgg16.VGG16(weights=None, include_top=False, input_shape=(32, 32, 3), seed=1)
    b = vgg16.VGG16(weights=None, include_top=False, input_shape=(32, 32, 3), seed=2)
    flat = parallel.pack_weights(a)
    assert flat.size == 14714688      # 13 conv kernels + biases
    parallel.unpack_weights(b, flat)
    for la, lb in zip(a.layers, b.layers):
        for wa, wb in zip(la.get_weights(), lb.get_weights()):
            assert np.array_equal(wa, wb)


def test_fasttape_equals_restatement():
    m = keras_shim.VGG16(include_top=False, input_shape=(64, 64, 3), seed=3, bias=True)
    img = synth.caffe_preprocess(synth.smooth_image_u8(0, 64, 64))
    ref = rs.visualize_all_layers(m, img, "block3_conv3", "all", only_layers=["block3_conv3"], top=4)
    entries = rs.forward(rs.build_entries(m, "block3_conv3"), img)
    feats = [f for f, _ in rs.find_top_filters(entries[-2].up, 4)]
    tape = rs.FastTape(m, "block3_conv3", img)
    got = tape.project(feats)
    for a, b in zip(got, ref["block3_conv3"]):
        assert rel_l2(a, b) < 5e-6      # conv_transpose2d vs conv2d(flipped): different fp32 summation order


WORKER = r'''
import os, sys
sys.path.insert(0, %r)
import numpy as np, torch, torch.distributed as dist
from deconv_api_b200 import parallel, vgg16
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
# every rank starts from DIFFERENT weights; after the one broadcast all hold rank 0's
m = vgg16.VGG16(weights=None, include_top=False, input_shape=(32, 32, 3), seed=100 + rank)
nbytes = parallel.broadcast_model_weights(m, dist, src=0)
ref = vgg16.VGG16(weights=None, include_top=False, input_shape=(32, 32, 3), seed=100)
for a, b in zip(m.layers, ref.layers):
    for wa, wb in zip(a.get_weights(), b.get_weights()):
        assert np.array_equal(wa, wb)
# image-sharded work: each rank owns distinct units, the union covers everything, nothing is exchanged
units = parallel.shard_work(6, 16, world, rank)
local = np.array([[i, lo, hi] for i, lo, hi in units], np.int64)
allu = parallel.gather_outputs(local, dist, dst=0)
if rank == 0:
    assert sorted(allu[:, 0].tolist()) == list(range(6))
# the sharded driver with a stub in place of the GPU engine: result[feature, image] encodes (image id, feature id).
# 5 images on 2 ranks -> image split (3 + 2, tapes of 2 images); 1 image on 2 ranks -> feature split, forward replicated
def runner(images, feats, out=None):
    n, B = len(feats), images.shape[0]
    r = np.zeros((n, B) + images.shape[1:], np.float32)
    for k, f in enumerate(feats):
        for b in range(B):
            r[k, b] = images[b] * 1000 + f
    return r
for N in (5, 1):
    imgs = np.arange(N, dtype=np.float32).reshape(N, 1, 1, 1) * np.ones((N, 2, 2, 3), np.float32)
    feats = [7, 3, 11, 2, 5]
    blocks = parallel.visualize_batch(m, imgs, "block1_conv1", features=feats, dist=dist, out_format="f32", images_per_tape=2, runner=runner)
    mine = sorted(i for idx, _, _ in blocks for i in idx)
    want_units = parallel.shard_work(N, len(feats), world, rank)
    assert mine == sorted(i for i, _, _ in want_units)
    for idx, fs, arr in blocks:
        assert arr.shape == (len(fs), len(idx), 2, 2, 3)
        assert list(fs) == feats[want_units[0][1]:want_units[0][2]]
    full = parallel.visualize_batch(m, imgs, "block1_conv1", features=feats, dist=dist, out_format="f32", images_per_tape=2, runner=runner, gather=True)
    if rank == 0:
        assert full.shape == (N, len(feats), 2, 2, 3)
        for i in range(N):
            for j, f in enumerate(feats):
                assert np.all(full[i, j] == i * 1000 + f), (N, i, f)
    else:
        assert full is None
dreams = parallel.dream_batch(m, np.arange(3, dtype=np.float32).reshape(3, 1, 1, 1) * np.ones((3, 2, 2, 3), np.float32), dist=dist, gather=True, runner=lambda img: img + 0.5)
if rank == 0:
    assert dreams.shape == (3, 2, 2, 3) and np.all(dreams[2] == 2.5)
    print("OK", nbytes)
dist.barrier()
dist.destroy_process_group()
'''


def test_gloo_world2_broadcast_and_sharding(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER % ROOT)
    env = dict(os.environ, OMP_NUM_THREADS="1")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29541", str(script)],
                         capture_output=True, text=True, timeout=300, env=env)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "OK 58" in out.stdout


def test_npz_weight_ingestion(tmp_path):
    """weights exported from Keras as <layer>/kernel (HWIO) and <layer>/bias load into the model."""
    src = vgg16.VGG16(weights=None, include_top=False, input_shape=(32, 32, 3), seed=5, bias=True)
    arrays = {}
    for l in src.layers:
        if l.kind == "conv":
            W, b = l.get_weights()
            arrays[l.name + "/kernel"], arrays[l.name + "/bias"] = W, b
    path = str(tmp_path / "vgg16_keras.npz")
    np.savez(path, **arrays)
    m = vgg16.VGG16(weights=path, include_top=False, input_shape=(32, 32, 3))
    for a, b in zip(src.layers, m.layers):
        for wa, wb in zip(a.get_weights(), b.get_weights()):
            assert np.array_equal(wa, wb)
    bad = dict(arrays)
    bad["block1_conv1/kernel"] = np.zeros((64, 3, 3, 3), np.float32)      # OIHW (torch layout) must be refused
    np.savez(path, **bad)
    import pytest
    with pytest.raises(ValueError):
        vgg16.VGG16(weights=path, include_top=False, input_shape=(32, 32, 3))


def test_imagenet_weights_never_fall_back_silently(tmp_path, monkeypatch):
    """weights='imagenet' without a weight file raises; DCV_ALLOW_SYNTH=1 turns that into a warning;
    an .npz export found through DCV_IMAGENET_NPZ is loaded."""
    monkeypatch.delenv("DCV_IMAGENET_NPZ", raising=False)
    monkeypatch.setenv("HOME", str(tmp_path))
    monkeypatch.setenv("DCV_ALLOW_SYNTH", "0")
    with pytest.raises(FileNotFoundError):
        vgg16.VGG16(weights="imagenet", include_top=False)
    monkeypatch.setenv("DCV_ALLOW_SYNTH", "1")
    with pytest.warns(RuntimeWarning, match="RANDOM"):
        m = vgg16.VGG16(weights="imagenet", include_top=False)
    arrays = {}
    for l in m.layers:
        if l.kind == "conv":
            W, b = l.get_weights()
            arrays[l.name + "/kernel"], arrays[l.name + "/bias"] = W * 0 + 0.25, b + 1.0
    path = tmp_path / "vgg.npz"
    np.savez(path, **arrays)
    monkeypatch.setenv("DCV_ALLOW_SYNTH", "0")
    monkeypatch.setenv("DCV_IMAGENET_NPZ", str(path))
    m2 = vgg16.VGG16(weights="imagenet", include_top=False)
    W, b = m2.get_layer("block2_conv1").get_weights()
    assert float(W.min()) == 0.25 and float(b.max()) == 1.0
