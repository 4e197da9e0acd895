"""Model build: the drop-in for ``keras.applications.vgg16`` as the reference uses it
(reference app/main.py:2,7,16-17: ``vgg16.VGG16(weights='imagenet', include_top=True)`` and
``preprocess_input``).

There is no network on the build/GPU boxes, so ``weights='imagenet'`` cannot download the Keras
``.h5``.  It resolves an ``.npz`` export of those weights instead (``DCV_IMAGENET_NPZ``, or
``~/.keras/models/vgg16_weights_tf_dim_ordering_tf_kernels.npz`` next to where Keras caches the
``.h5``) and RAISES when none exists: a drop-in user must never get a random network silently.
``DCV_ALLOW_SYNTH=1`` downgrades that error to a warning and falls back to the seeded init.
``weights=None`` gives the seeded Glorot-uniform init of ``synth`` (Keras' own default initialiser),
and ``weights='<file>.npz'`` loads arrays saved under the Keras names ``<layer>/kernel`` (HWIO /
(in,out)) and ``<layer>/bias``.
"""
import os
import warnings

import numpy as np

from . import synth
from .layers import (Convolution2D, Dense, Flatten, InputLayer, MaxPooling2D, Model)


IMAGENET_NPZ_DEFAULT = os.path.join("~", ".keras", "models", "vgg16_weights_tf_dim_ordering_tf_kernels.npz")


def imagenet_npz_path():
    """Path of an .npz export of the Keras ImageNet VGG16 weights, or None."""
    for cand in (os.environ.get("DCV_IMAGENET_NPZ"), os.path.expanduser(IMAGENET_NPZ_DEFAULT)):
        if cand and os.path.isfile(cand):
            return cand
    return None


def VGG16(include_top=True, weights="imagenet", input_shape=None, seed=1234, bias=False):
    if input_shape is None:
        input_shape = (224, 224, 3) if include_top else (None, None, 3)
    npz = None
    if isinstance(weights, str) and weights.endswith(".npz"):
        npz = np.load(weights)
    elif weights == "imagenet":
        path = imagenet_npz_path()
        if path is not None:
            npz = np.load(path)
        elif os.environ.get("DCV_ALLOW_SYNTH", "0") == "1":
            warnings.warn("VGG16(weights='imagenet'): no ImageNet weight file found (DCV_IMAGENET_NPZ, %s); "
                          "DCV_ALLOW_SYNTH=1 -> using seeded RANDOM weights, the visualisations are of an untrained network"
                          % IMAGENET_NPZ_DEFAULT, RuntimeWarning, stacklevel=2)
        else:
            raise FileNotFoundError(
                "VGG16(weights='imagenet'): the Keras ImageNet weights cannot be downloaded here (no network) and no "
                ".npz export was found; set DCV_IMAGENET_NPZ (or DCV_WEIGHTS for the service) to a file with the arrays "
                "'<layer>/kernel' and '<layer>/bias', put it at %s, pass weights=None for the seeded random init, "
                "or set DCV_ALLOW_SYNTH=1 to accept random weights with a warning" % IMAGENET_NPZ_DEFAULT)
    elif weights is not None:
        raise ValueError("weights must be None, 'imagenet' or a path to an .npz file")
    conv_w = dict(zip([n for n, _, _ in synth.VGG16_CONVS], synth.conv_weights(seed, bias)))
    layers = []
    for kind, name in synth.VGG16_TOPOLOGY:
        if kind == "input":
            layers.append(InputLayer(input_shape, name=name))
        elif kind == "conv":
            W, b = conv_w[name]
            if npz is not None:
                W, b = npz[name + "/kernel"], npz[name + "/bias"]
                if W.shape != conv_w[name][0].shape:
                    raise ValueError("%s/kernel has shape %s, expected HWIO %s" % (name, W.shape, conv_w[name][0].shape))
            l = Convolution2D(W.shape[3], (3, 3), activation="relu", padding="same", name=name)
            l._weights = [np.asarray(W, np.float32), np.asarray(b, np.float32)]
            layers.append(l)
        elif kind == "pool":
            layers.append(MaxPooling2D((2, 2), strides=(2, 2), name=name))
        elif not include_top:
            continue
        elif kind == "flatten":
            layers.append(Flatten(name=name))
        elif kind == "dense":
            units = dict((n, o) for n, _, o in synth.VGG16_DENSE)[name]
            act = "softmax" if name == "predictions" else "relu"
            if npz is not None:
                lazy = (lambda n=name: (npz[n + "/kernel"], npz[n + "/bias"]))
            else:
                lazy = (lambda n=name: synth.dense_weights(n, seed, bias))
            layers.append(Dense(units, activation=act, name=name, lazy=lazy))
    return Model(layers, name="vgg16")


def preprocess_input(x, data_format=None, mode="caffe"):
    """Keras-Applications 1.0.8 'caffe' mode (what reference main.py:61 applies): reverse the channel
    axis and subtract [103.939, 116.779, 123.68].  Pure data preparation on a 150 KB image; the GPU
    path starts at ``visualize_all_layers``."""
    if mode != "caffe":
        raise NotImplementedError("only the 'caffe' mode that VGG16 uses is provided")
    x = np.asarray(x)
    if not np.issubdtype(x.dtype, np.floating):
        x = x.astype(np.float32)
    x = x[..., ::-1]
    x = x - synth.CAFFE_MEAN.astype(x.dtype)
    return x
