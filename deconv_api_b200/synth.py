"""Seeded synthetic weights and images (SURVEY.md §8d "Synthetic inputs").

There is no network on the build or GPU boxes, so the ImageNet ``.h5`` that the
reference downloads at import (reference app/main.py:17) is unavailable; every
run uses Glorot-uniform random weights instead.  The generators here are the
single source of truth for both the CUDA path and the oracle so that the two
always see bit-identical inputs.
"""
import numpy as np

# (name, Cin, Cout) of the 13 VGG16 conv layers in topology order
# (Keras-Applications 1.0.8 layer names; reference app/requirements.txt:55).
VGG16_CONVS = [
    ("block1_conv1", 3, 64), ("block1_conv2", 64, 64),
    ("block2_conv1", 64, 128), ("block2_conv2", 128, 128),
    ("block3_conv1", 128, 256), ("block3_conv2", 256, 256), ("block3_conv3", 256, 256),
    ("block4_conv1", 256, 512), ("block4_conv2", 512, 512), ("block4_conv3", 512, 512),
    ("block5_conv1", 512, 512), ("block5_conv2", 512, 512), ("block5_conv3", 512, 512),
]
# (name, units_in, units_out) of the classifier head (include_top=True).
VGG16_DENSE = [("fc1", 25088, 4096), ("fc2", 4096, 4096), ("predictions", 4096, 1000)]

# Full layer order of the Keras VGG16 graph: ("conv", name) / ("pool", name) / ...
VGG16_TOPOLOGY = (
    [("input", "input_1")]
    + [("conv", "block1_conv1"), ("conv", "block1_conv2"), ("pool", "block1_pool")]
    + [("conv", "block2_conv1"), ("conv", "block2_conv2"), ("pool", "block2_pool")]
    + [("conv", "block3_conv1"), ("conv", "block3_conv2"), ("conv", "block3_conv3"), ("pool", "block3_pool")]
    + [("conv", "block4_conv1"), ("conv", "block4_conv2"), ("conv", "block4_conv3"), ("pool", "block4_pool")]
    + [("conv", "block5_conv1"), ("conv", "block5_conv2"), ("conv", "block5_conv3"), ("pool", "block5_pool")]
    + [("flatten", "flatten"), ("dense", "fc1"), ("dense", "fc2"), ("dense", "predictions")]
)

CAFFE_MEAN = np.array([103.939, 116.779, 123.68], dtype=np.float32)


def glorot_conv(rng, cin, cout, k=3):
    """Keras ``glorot_uniform`` for a (k,k,cin,cout) HWIO kernel: limit = sqrt(6/(fan_in+fan_out))."""
    limit = np.sqrt(6.0 / (k * k * cin + k * k * cout))
    return rng.uniform(-limit, limit, (k, k, cin, cout)).astype(np.float32)


def conv_weights(seed=1234, bias=False, convs=VGG16_CONVS):
    """List of (W_hwio f32, b f32) per conv layer, drawn in topology order from one rng stream.

    bias=False gives the Keras default (zeros); bias=True draws b ~ U(-0.1, 0.1) *after* all the
    kernels so that the kernels are identical in both settings.
    """
    rng = np.random.default_rng(seed)
    ws = [glorot_conv(rng, cin, cout) for _, cin, cout in convs]
    if bias:
        bs = [rng.uniform(-0.1, 0.1, (cout,)).astype(np.float32) for _, _, cout in convs]
    else:
        bs = [np.zeros((cout,), np.float32) for _, _, cout in convs]
    return list(zip(ws, bs))


def dense_weights(name, seed=1234, bias=False):
    """(W (in,out) f32, b) for one classifier layer; its own rng stream so it can be built lazily
    (fc1 alone is 411 MB)."""
    idx = [n for n, _, _ in VGG16_DENSE].index(name)
    _, nin, nout = VGG16_DENSE[idx]
    rng = np.random.default_rng([seed, 1000 + idx])
    limit = np.sqrt(6.0 / (nin + nout))
    w = rng.uniform(-limit, limit, (nin, nout)).astype(np.float32)
    b = (rng.uniform(-0.1, 0.1, (nout,)).astype(np.float32) if bias else np.zeros((nout,), np.float32))
    return w, b


def image_u8(seed, h=224, w=224):
    """Uniform-random uint8 HxWx3 image (stands for what cv2.imdecode returns: BGR)."""
    return np.random.default_rng(seed).integers(0, 256, (h, w, 3), dtype=np.uint8)


def smooth_image_u8(seed, h=224, w=224):
    """A low-frequency synthetic image (sum of a few sinusoids + mild noise).  Natural-image-like
    statistics keep deep activations alive better than white noise."""
    rng = np.random.default_rng(seed)
    yy, xx = np.meshgrid(np.linspace(0, 1, h, dtype=np.float32), np.linspace(0, 1, w, dtype=np.float32), indexing="ij")
    img = np.zeros((h, w, 3), np.float32)
    for c in range(3):
        for _ in range(6):
            fx, fy = rng.uniform(0.5, 9.0, 2)
            ph = rng.uniform(0, 2 * np.pi)
            img[..., c] += rng.uniform(0.3, 1.0) * np.sin(2 * np.pi * (fx * xx + fy * yy) + ph)
    img = (img - img.min()) / (img.max() - img.min())
    img = img * 255.0 + rng.normal(0, 4.0, img.shape)
    return np.clip(img, 0, 255).astype(np.uint8)


def caffe_preprocess(img_u8):
    """What reference main.py:60-61 does to a decoded image: float32, reverse the channel axis,
    subtract the per-channel means.  Returns (1,H,W,3) float32."""
    x = img_u8.astype(np.float32)[..., ::-1]
    x = x - CAFFE_MEAN
    return np.ascontiguousarray(x[None])


def image_batch(n, h=224, w=224, seed0=0):
    return np.concatenate([caffe_preprocess(image_u8(seed0 + i, h, w)) for i in range(n)], axis=0)
