"""The sharded product path (parallel.visualize_batch / dream_batch, BASELINE configs[3] / [4]) on the GPU: one process
against the oracle, and two gloo ranks sharing cuda:0 for the feature-split branch (fewer images than ranks: every
rank replicates the forward and projects a contiguous slice of the feature maps, results gathered on rank 0)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from oracle import keras_shim, restatement as rs
from deconv_api_b200 import dream as dm, parallel, synth, vgg16
from conftest import rel_l2

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_visualize_batch_one_process_vs_oracle():
    model = vgg16.VGG16(weights=None, include_top=False, input_shape=(64, 64, 3), seed=11, bias=True)
    ref_model = keras_shim.VGG16(include_top=False, input_shape=(64, 64, 3), seed=11, bias=True)
    imgs = np.concatenate([synth.caffe_preprocess(synth.smooth_image_u8(s, 64, 64)) for s in (1, 2, 3)], axis=0)
    feats = [0, 5, 9, 17, 33, 64, 100, 127, 3, 77]
    full = parallel.visualize_batch(model, imgs, "block2_conv2", features=feats, precision="fp32", out_format="f32",
                                    images_per_tape=2, gather=True)
    assert full.shape == (3, 10, 64, 64, 3)
    want = np.stack([rs.FastTape(ref_model, "block2_conv2", imgs[i:i + 1]).project(feats) for i in range(3)])
    for i in range(3):
        for j in range(10):
            if np.abs(want[i, j]).max() > 0:
                assert rel_l2(full[i, j], want[i, j]) < 1e-5, (i, j)
    # default product formats: bf16 chain, uint8 results left sharded as blocks (feature-major inside a block)
    blocks = parallel.visualize_batch(model, imgs, "block2_conv2", features=feats, images_per_tape=2)
    assert [idx for idx, _, _ in blocks] == [[0, 1], [2]]
    for idx, fs, arr in blocks:
        assert arr.dtype == np.uint8 and arr.shape == (10, len(idx), 64, 64, 3)
        for b, i in enumerate(idx):
            for j in range(10):
                ref8 = rs.deprocess_image(want[i, j].copy())
                d = np.abs(arr[j, b].astype(np.int16) - ref8.astype(np.int16))
                assert np.percentile(d, 99) <= 2 and d.mean() < 0.5, (i, j, d.max())      # bf16 chain (2e-2) then 8-bit levels


WORKER = r'''
import os, sys
sys.path.insert(0, %r)
os.environ["DCV_DEVICE"] = "0"
import numpy as np, torch, torch.distributed as dist
from deconv_api_b200 import parallel, synth, vgg16
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
model = vgg16.VGG16(weights=None, include_top=False, input_shape=(64, 64, 3), seed=100 + rank)
parallel.broadcast_model_weights(model, dist, src=0)
img = synth.caffe_preprocess(synth.smooth_image_u8(4, 64, 64))
feats = list(range(0, 128, 8))
full = parallel.visualize_batch(model, img, "block2_conv2", features=feats, dist=dist, precision="fp32", out_format="f32", gather=True)
blocks = parallel.visualize_batch(model, img, "block2_conv2", features=feats, dist=dist, precision="fp32", out_format="f32")
assert len(blocks) == 1 and list(blocks[0][1]) == feats[rank * 8:(rank + 1) * 8]       # feature split, forward replicated
if rank == 0:
    np.save(sys.argv[1], full)
    print("OK")
dist.barrier()
dist.destroy_process_group()
'''


def test_two_ranks_feature_split_on_one_gpu(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER % ROOT)
    out_npy = str(tmp_path / "full.npy")
    env = dict(os.environ, OMP_NUM_THREADS="1")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29547", str(script), out_npy],
                         capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0 and "OK" in out.stdout, out.stdout + out.stderr
    full = np.load(out_npy)
    assert full.shape == (1, 16, 64, 64, 3)
    ref_model = keras_shim.VGG16(include_top=False, input_shape=(64, 64, 3), seed=100)
    img = synth.caffe_preprocess(synth.smooth_image_u8(4, 64, 64))
    want = rs.FastTape(ref_model, "block2_conv2", img).project(list(range(0, 128, 8)))
    for j in range(16):
        if np.abs(want[j]).max() > 0:
            assert rel_l2(full[0, j], want[j]) < 1e-5, j


def test_dream_batch_matches_single_dreams():
    model = vgg16.VGG16(weights=None, include_top=False, input_shape=(None, None, 3), seed=9)
    imgs = np.concatenate([synth.caffe_preprocess(synth.smooth_image_u8(s, 96, 80)) for s in (4, 5)], axis=0)
    kw = dict(layers={"block2_conv2": 0.5, "block3_conv3": 2.0}, num_octave=2, iterations=3, jitter=8, seed=3, precision="bf16")
    got = parallel.dream_batch(model, imgs, gather=True, **kw)
    assert got.shape == imgs.shape
    for i in range(2):
        assert np.array_equal(got[i], dm.dream(model, imgs[i], **kw))
