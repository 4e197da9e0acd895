"""TEST INFRASTRUCTURE ONLY — a minimal stand-in for Keras 2.3.1 (reference app/requirements.txt:54).

TensorFlow/Keras cannot be installed here (no network).  This module fakes exactly the Keras
surface that the reference's ``app/deepdream.py`` and ``app/main.py`` touch (list in SURVEY.md
§8c) so that those two files run UNMODIFIED; conv/dense arithmetic is torch-CPU fp32
(cross-correlation, ``padding='same'``, stride 1, NHWC/HWIO — the Keras channels_last contract).
Everything else on the path (pool-with-switch, unpool, top-k, masking, de-process) is then the
reference's own numpy code, which is what makes it the oracle.

``install()`` registers the fake package in ``sys.modules`` under the name ``keras``.
Nothing in the product package may import this file.
"""
import sys
import types

import numpy as np
import torch
import torch.nn.functional as F

_STATE = {"fmt": "channels_last", "uid": {}}


def _uid(prefix):
    n = _STATE["uid"].get(prefix, 0) + 1
    _STATE["uid"][prefix] = n
    return "%s_%d" % (prefix, n)


def reset_uids():
    _STATE["uid"].clear()


# ----------------------------------------------------------------------------- symbolic tensors
class Sym(object):
    """Lazy tensor: ``fn(env) -> ndarray`` where env maps placeholder id -> fed value."""

    def __init__(self, fn, shape, layer=None, parent=None):
        self.fn = fn
        self._keras_shape = tuple(shape)
        self.layer = layer      # producing layer (None for raw placeholders / backend ops)
        self.parent = parent    # input Sym of the producing layer

    @property
    def shape(self):
        return self._keras_shape

    def eval(self, env):
        # a fed tensor cuts the graph (K.function([layer.input], [layer.output]) feeds mid-graph)
        if id(self) in env:
            return env[id(self)]
        return self.fn(env)


def _placeholder(shape):
    s = Sym(None, shape)
    s.fn = lambda env, _s=s: env[id(_s)]
    return s


_LEARNING_PHASE = _placeholder(())


def _evaluate(outputs, inputs, values):
    env = {}
    for ph, v in zip(inputs, values):
        env[id(ph)] = v
    return [np.asarray(o.eval(env)) for o in outputs]


# ----------------------------------------------------------------------------- activations
def _relu(x):
    if isinstance(x, Sym):
        return Sym(lambda env, _x=x: np.maximum(_x.eval(env), 0).astype(np.float32), x.shape)
    return np.maximum(x, 0)


def _softmax(x):
    def f(v):
        v = np.asarray(v, np.float32)
        e = np.exp(v - v.max(axis=-1, keepdims=True))
        return (e / e.sum(axis=-1, keepdims=True)).astype(np.float32)
    if isinstance(x, Sym):
        return Sym(lambda env, _x=x: f(_x.eval(env)), x.shape)
    return f(x)


def _linear(x):
    return x


_relu.__name__ = "relu"
_softmax.__name__ = "softmax"
_linear.__name__ = "linear"
_ACT = {"relu": _relu, "softmax": _softmax, "linear": _linear, None: _linear}


# ----------------------------------------------------------------------------- layers
class Layer(object):
    def __init__(self, name=None, **_):
        self.name = name or _uid(type(self).__name__.lower())
        self.input = None
        self.output = None
        self.input_shape = None
        self.output_shape = None
        self._weights = None

    # -- graph building
    def __call__(self, x):
        self.input = x
        self.input_shape = tuple(x.shape)
        self.output_shape = tuple(self.compute_output_shape(self.input_shape))
        self.build(self.input_shape)
        out = Sym(lambda env, _x=x: self.call(_x.eval(env)), self.output_shape, layer=self, parent=x)
        self.output = out
        return out

    def build(self, input_shape):
        pass

    def compute_output_shape(self, s):
        return s

    def get_weights(self):
        return [w.copy() for w in (self._weights or [])]

    def set_weights(self, ws):
        ws = [np.asarray(w, dtype=np.float32) for w in ws]
        if self._weights is not None and len(self._weights):
            assert len(ws) == len(self._weights), "weight count mismatch"
            for a, b in zip(ws, self._weights):
                assert a.shape == b.shape, "weight shape mismatch %s vs %s" % (a.shape, b.shape)
        self._weights = ws

    def get_config(self):
        return {"name": self.name}

    @classmethod
    def from_config(cls, config):
        return cls(**config)


class InputLayer(Layer):
    def __init__(self, input_shape=None, name=None, **_):
        super(InputLayer, self).__init__(name=name or _uid("input"))
        shape = (None,) + tuple(input_shape)
        self.input_shape = shape
        self.output_shape = shape
        ph = _placeholder(shape)
        ph.layer = self
        self.input = ph
        self.output = ph

    def get_config(self):
        return {"name": self.name, "batch_input_shape": self.input_shape}


def Input(shape=None, name=None, **_):
    return InputLayer(input_shape=shape, name=name).output


class Conv2D(Layer):
    def __init__(self, filters, kernel_size, strides=(1, 1), padding="valid", activation=None,
                 use_bias=True, name=None, **_):
        super(Conv2D, self).__init__(name=name or _uid("conv2d"))
        self.filters = int(filters)
        ks = kernel_size if isinstance(kernel_size, (tuple, list)) else (kernel_size, kernel_size)
        self.kernel_size = (int(ks[0]), int(ks[1]))
        st = strides if isinstance(strides, (tuple, list)) else (strides, strides)
        self.strides = (int(st[0]), int(st[1]))
        assert self.strides == (1, 1), "shim supports stride 1 only (all VGG16 convs)"
        self.padding = padding
        self._act_name = activation if isinstance(activation, (str, type(None))) else activation.__name__
        self.activation = _ACT[self._act_name]
        self.use_bias = use_bias

    def build(self, input_shape):
        if self._weights is None:
            cin = input_shape[-1]
            self._weights = [np.zeros(self.kernel_size + (cin, self.filters), np.float32),
                             np.zeros((self.filters,), np.float32)]

    def compute_output_shape(self, s):
        if self.padding == "same":
            return (s[0], s[1], s[2], self.filters)
        kh, kw = self.kernel_size
        return (s[0], None if s[1] is None else s[1] - kh + 1, None if s[2] is None else s[2] - kw + 1, self.filters)

    def call(self, x):
        W, b = self._weights
        xt = torch.from_numpy(np.ascontiguousarray(np.asarray(x, np.float32))).permute(0, 3, 1, 2)
        wt = torch.from_numpy(np.ascontiguousarray(W)).permute(3, 2, 0, 1).contiguous()
        kh, kw = self.kernel_size
        pad = (kh // 2, kw // 2) if self.padding == "same" else 0
        y = F.conv2d(xt, wt, torch.from_numpy(b) if self.use_bias else None, padding=pad)
        y = y.permute(0, 2, 3, 1).contiguous().numpy()
        return self.activation(y)

    def get_config(self):
        return {"name": self.name, "filters": self.filters, "kernel_size": self.kernel_size,
                "strides": self.strides, "padding": self.padding, "activation": self._act_name,
                "use_bias": self.use_bias}


Convolution2D = Conv2D


class MaxPooling2D(Layer):
    def __init__(self, pool_size=(2, 2), strides=None, padding="valid", name=None, **_):
        super(MaxPooling2D, self).__init__(name=name or _uid("max_pooling2d"))
        self.pool_size = tuple(pool_size)
        self.strides = tuple(strides) if strides else self.pool_size
        self.padding = padding

    def compute_output_shape(self, s):
        return (s[0], None if s[1] is None else s[1] // self.pool_size[0], None if s[2] is None else s[2] // self.pool_size[1], s[3])

    def call(self, x):
        xt = torch.from_numpy(np.ascontiguousarray(np.asarray(x, np.float32))).permute(0, 3, 1, 2)
        return F.max_pool2d(xt, self.pool_size, self.strides).permute(0, 2, 3, 1).contiguous().numpy()

    def get_config(self):
        return {"name": self.name, "pool_size": self.pool_size, "strides": self.strides, "padding": self.padding}


class Flatten(Layer):
    def compute_output_shape(self, s):
        return (s[0], int(np.prod(s[1:])))

    def call(self, x):
        x = np.asarray(x)
        return x.reshape(x.shape[0], -1)


class Activation(Layer):
    def __init__(self, activation, name=None, **_):
        super(Activation, self).__init__(name=name or _uid("activation"))
        self._act_name = activation
        self.activation = _ACT[activation]

    def call(self, x):
        return self.activation(np.asarray(x, np.float32))

    def get_config(self):
        return {"name": self.name, "activation": self._act_name}


class Dense(Layer):
    def __init__(self, units, activation=None, use_bias=True, name=None, _lazy=None, **_):
        super(Dense, self).__init__(name=name or _uid("dense"))
        self.units = int(units)
        self._act_name = activation if isinstance(activation, (str, type(None))) else activation.__name__
        self.activation = _ACT[self._act_name]
        self.use_bias = use_bias
        self._lazy = _lazy      # callable -> [W, b]; lets the 411 MB fc1 matrix exist only when used

    def build(self, input_shape):
        if self._weights is None and self._lazy is None:
            self._weights = [np.zeros((input_shape[-1], self.units), np.float32),
                             np.zeros((self.units,), np.float32)]

    def compute_output_shape(self, s):
        return (s[0], self.units)

    def _materialise(self):
        if self._weights is None and self._lazy is not None:
            self._weights = [np.asarray(w, np.float32) for w in self._lazy()]

    def get_weights(self):
        self._materialise()
        return super(Dense, self).get_weights()

    def call(self, x):
        self._materialise()
        W, b = self._weights
        y = torch.from_numpy(np.ascontiguousarray(np.asarray(x, np.float32))) @ torch.from_numpy(W)
        if self.use_bias:
            y = y + torch.from_numpy(b)
        return self.activation(y.numpy())

    def get_config(self):
        return {"name": self.name, "units": self.units, "activation": self._act_name, "use_bias": self.use_bias}


# ----------------------------------------------------------------------------- Model
class Model(object):
    def __init__(self, input, output, name=None):
        self.input = input
        self.output = output
        chain = []
        t = output
        while t is not None and t.layer is not None:
            chain.append(t.layer)
            t = t.parent
        self.layers = list(reversed(chain))
        self.name = name or _uid("model")

    def predict(self, x, **_):
        return _evaluate([self.output], [self.input], [np.asarray(x, np.float32)])[0].astype(np.float32)

    def set_weights(self, ws):
        ws = list(ws)
        for l in self.layers:
            n = len(l._weights) if l._weights else 0
            if n:
                l.set_weights(ws[:n])
                ws = ws[n:]
        assert not ws, "too many weights"

    def get_weights(self):
        out = []
        for l in self.layers:
            out += l.get_weights()
        return out


# ----------------------------------------------------------------------------- backend
def _k_function(inputs, outputs):
    def run(values):
        return _evaluate(outputs, inputs, values)
    return run


# ----------------------------------------------------------------------------- applications.vgg16
def VGG16(include_top=True, weights="imagenet", input_shape=None, seed=1234, bias=False, **_):
    """VGG16 graph with the Keras-Applications 1.0.8 layer names.  ``weights`` is accepted for
    signature compatibility; ImageNet weights cannot be downloaded here, so every value of it gives
    the seeded Glorot-uniform init of ``deconv_api_b200.synth`` (shared with the CUDA path)."""
    from deconv_api_b200 import synth
    reset_uids()
    if input_shape is None:
        input_shape = (224, 224, 3) if include_top else (None, None, 3)
    x = Input(shape=input_shape, name="input_1")
    inp = x
    conv_w = dict(zip([n for n, _, _ in synth.VGG16_CONVS], synth.conv_weights(seed, bias)))
    for kind, name in synth.VGG16_TOPOLOGY:
        if kind == "conv":
            cout = conv_w[name][0].shape[3]
            l = Conv2D(cout, (3, 3), activation="relu", padding="same", name=name)
            x = l(x)
            l.set_weights(conv_w[name])
        elif kind == "pool":
            x = MaxPooling2D((2, 2), strides=(2, 2), name=name)(x)
        elif not include_top:
            continue
        elif kind == "flatten":
            x = Flatten(name=name)(x)
        elif kind == "dense":
            units = dict((n, o) for n, _, o in synth.VGG16_DENSE)[name]
            act = "softmax" if name == "predictions" else "relu"
            x = Dense(units, activation=act, name=name,
                      _lazy=(lambda n=name: synth.dense_weights(n, seed, bias)))(x)
    return Model(inp, x, name="vgg16")


def preprocess_input(x, data_format=None, mode="caffe"):
    """'caffe' mode of Keras-Applications 1.0.8 imagenet_utils: RGB->BGR flip, subtract mean.
    Mutates a float ndarray argument in place like the original."""
    assert mode == "caffe"
    if not np.issubdtype(x.dtype, np.floating):
        x = x.astype(np.float32)
    x = x[..., ::-1]
    mean = [103.939, 116.779, 123.68]
    x[..., 0] -= mean[0]
    x[..., 1] -= mean[1]
    x[..., 2] -= mean[2]
    return x


def img_to_array(img, data_format=None, dtype="float32"):
    x = np.asarray(img, dtype=dtype)
    if x.ndim == 2:
        x = x[..., None]
    return x


# ----------------------------------------------------------------------------- install
def install():
    """Register the fake ``keras`` package tree in sys.modules (idempotent)."""
    if "keras" in sys.modules and getattr(sys.modules["keras"], "__dcv_shim__", False):
        return sys.modules["keras"]

    def mod(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    conv = mod("keras.layers.convolutional", Convolution2D=Conv2D, Conv2D=Conv2D, MaxPooling2D=MaxPooling2D)
    layers = mod("keras.layers", Input=Input, InputLayer=InputLayer, Flatten=Flatten, Activation=Activation,
                 Dense=Dense, Conv2D=Conv2D, Convolution2D=Conv2D, MaxPooling2D=MaxPooling2D, convolutional=conv)
    models = mod("keras.models", Model=Model)
    vgg = mod("keras.applications.vgg16", VGG16=VGG16, preprocess_input=preprocess_input)
    apps = mod("keras.applications", vgg16=vgg)

    class _Scope(object):
        value = True

    tfb = mod("keras.backend.tensorflow_backend", _SYMBOLIC_SCOPE=_Scope())

    def set_image_data_format(f):
        _STATE["fmt"] = f

    backend = mod("keras.backend", set_image_data_format=set_image_data_format,
                  image_data_format=lambda: _STATE["fmt"], epsilon=lambda: 1e-7,
                  placeholder=lambda shape=None, **_: _placeholder(shape),
                  learning_phase=lambda: _LEARNING_PHASE, function=_k_function,
                  tensorflow_backend=tfb)
    image = mod("keras.preprocessing.image", img_to_array=img_to_array)
    prep = mod("keras.preprocessing", image=image)
    keras = mod("keras", layers=layers, models=models, applications=apps, backend=backend,
                preprocessing=prep, __dcv_shim__=True, __version__="2.3.1-shim")
    return keras


# ----------------------------------------------------------------------------- small test model
def mini_model(seed=7, hw=16, chans=(8, 8, 16), dense=(32, 10), bias=True):
    """A tiny VGG-shaped graph (conv,conv,pool,conv,pool,flatten,dense,dense-softmax) so that the
    whole reference driver — dense/flatten targets included — runs in milliseconds and its full
    output fits in a committed fixture."""
    rng = np.random.default_rng(seed)
    reset_uids()

    def cw(cin, cout):
        lim = np.sqrt(6.0 / (9 * cin + 9 * cout))
        return [rng.uniform(-lim, lim, (3, 3, cin, cout)).astype(np.float32),
                (rng.uniform(-0.1, 0.1, (cout,)) if bias else np.zeros((cout,))).astype(np.float32)]

    def dw(nin, nout):
        lim = np.sqrt(6.0 / (nin + nout))
        return [rng.uniform(-lim, lim, (nin, nout)).astype(np.float32),
                (rng.uniform(-0.1, 0.1, (nout,)) if bias else np.zeros((nout,))).astype(np.float32)]

    x = inp = Input(shape=(hw, hw, 3), name="input_1")
    plan = [("conv", "block1_conv1", chans[0]), ("conv", "block1_conv2", chans[1]), ("pool", "block1_pool", 0),
            ("conv", "block2_conv1", chans[2]), ("pool", "block2_pool", 0)]
    cin = 3
    for kind, name, cout in plan:
        if kind == "conv":
            l = Conv2D(cout, (3, 3), activation="relu", padding="same", name=name)
            x = l(x)
            l.set_weights(cw(cin, cout))
            cin = cout
        else:
            x = MaxPooling2D((2, 2), strides=(2, 2), name=name)(x)
    x = Flatten(name="flatten")(x)
    nin = (hw // 4) * (hw // 4) * cin
    l = Dense(dense[0], activation="relu", name="fc1")
    x = l(x)
    l.set_weights(dw(nin, dense[0]))
    l = Dense(dense[1], activation="softmax", name="predictions")
    x = l(x)
    l.set_weights(dw(dense[0], dense[1]))
    return Model(inp, x, name="mini")
