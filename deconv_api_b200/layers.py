"""Keras-shaped layer / model description objects.

The reference drives Keras objects (reference app/deepdream.py:403-417 does ``isinstance`` checks
against Convolution2D / MaxPooling2D / Dense / Activation / Flatten / InputLayer and reads
``name, get_weights(), get_config(), input_shape, output_shape, activation, pool_size``).  These
classes expose the same attributes so that a model built here can be handed to the drop-in
``visualize_all_layers``; they hold weights on the host and describe the graph — the arithmetic
happens in the CUDA library once the model is bound to an Engine.
"""
import numpy as np

from . import engine as _engine


class Layer(object):
    kind = None

    def __init__(self, name):
        self.name = name
        self.input_shape = None
        self.output_shape = None
        self._weights = []

    def get_weights(self):
        return [w.copy() for w in self._weights]

    def set_weights(self, ws):
        ws = [np.asarray(w, dtype=np.float32) for w in ws]
        if self._weights:
            assert len(ws) == len(self._weights)
            for a, b in zip(ws, self._weights):
                assert a.shape == b.shape, "weight shape mismatch %s vs %s" % (a.shape, b.shape)
        self._weights = ws
        self._version = getattr(self, "_version", 0) + 1

    def get_config(self):
        return {"name": self.name}

    def _build(self, input_shape):
        self.input_shape = tuple(input_shape)
        self.output_shape = tuple(self.compute_output_shape(self.input_shape))

    def compute_output_shape(self, s):
        return s


class InputLayer(Layer):
    kind = "input"

    def __init__(self, input_shape, name="input_1"):
        Layer.__init__(self, name)
        self.input_shape = self.output_shape = (None,) + tuple(input_shape)

    def get_config(self):
        return {"name": self.name, "batch_input_shape": self.input_shape}


class Convolution2D(Layer):
    kind = "conv"

    def __init__(self, filters, kernel_size=(3, 3), activation=None, padding="same", use_bias=True, name=None):
        Layer.__init__(self, name)
        self.filters = int(filters)
        self.kernel_size = tuple(kernel_size)
        if self.kernel_size != (3, 3) or padding != "same":
            raise NotImplementedError("the CUDA path implements 3x3 / stride 1 / padding='same' convolutions (all of VGG16)")
        self.padding = padding
        self.strides = (1, 1)
        self.use_bias = use_bias
        self.activation_name = activation
        self.activation = activation      # the reference only passes it on to DActivation

    def _build(self, input_shape):
        Layer._build(self, input_shape)
        if not self._weights:
            self._weights = [np.zeros(self.kernel_size + (input_shape[-1], self.filters), np.float32),
                             np.zeros((self.filters,), np.float32)]

    def compute_output_shape(self, s):
        return (s[0], s[1], s[2], self.filters)

    def get_config(self):
        return {"name": self.name, "filters": self.filters, "kernel_size": self.kernel_size, "strides": self.strides,
                "padding": self.padding, "activation": self.activation_name, "use_bias": self.use_bias}


Conv2D = Convolution2D


class MaxPooling2D(Layer):
    kind = "pool"

    def __init__(self, pool_size=(2, 2), strides=None, name=None):
        Layer.__init__(self, name)
        self.pool_size = tuple(pool_size)
        self.strides = tuple(strides) if strides else self.pool_size
        if self.pool_size != (2, 2) or self.strides != (2, 2):
            raise NotImplementedError("the CUDA path implements 2x2 / stride 2 max pooling (all of VGG16)")

    def compute_output_shape(self, s):
        return (s[0], None if s[1] is None else s[1] // 2, None if s[2] is None else s[2] // 2, s[3])

    def get_config(self):
        return {"name": self.name, "pool_size": self.pool_size, "strides": self.strides, "padding": "valid"}


class Flatten(Layer):
    kind = "flatten"

    def compute_output_shape(self, s):
        return (s[0], int(np.prod(s[1:])))


class Dense(Layer):
    kind = "dense"

    def __init__(self, units, activation=None, use_bias=True, name=None, lazy=None):
        Layer.__init__(self, name)
        self.units = int(units)
        self.activation_name = activation
        self.activation = activation
        self.use_bias = use_bias
        self._lazy = lazy

    def _build(self, input_shape):
        Layer._build(self, input_shape)
        if not self._weights and self._lazy is None:
            self._weights = [np.zeros((input_shape[-1], self.units), np.float32), np.zeros((self.units,), np.float32)]

    def get_weights(self):
        if not self._weights and self._lazy is not None:
            self._weights = [np.asarray(w, np.float32) for w in self._lazy()]
        return Layer.get_weights(self)

    def compute_output_shape(self, s):
        return (s[0], self.units)

    def get_config(self):
        return {"name": self.name, "units": self.units, "activation": self.activation_name, "use_bias": self.use_bias}


class Activation(Layer):
    kind = "activation"

    def __init__(self, activation, name=None):
        Layer.__init__(self, name)
        self.activation_name = activation
        self.activation = activation


class Model(object):
    """A sequential stack of the layers above (VGG16 is sequential)."""

    def __init__(self, layers, name="model"):
        self.layers = list(layers)
        self.name = name
        shape = self.layers[0].output_shape
        for l in self.layers[1:]:
            l._build(shape)
            shape = l.output_shape
        self.input_shape = self.layers[0].input_shape
        self.output_shape = shape
        self._engines = {}

    def get_layer(self, name):
        for l in self.layers:
            if l.name == name:
                return l
        raise ValueError("No such layer: " + name)

    def bound_engine(self, device=None, upto=None):
        """Engine with this model's network loaded through layer `upto` (name; None = all).  Layers
        are uploaded once and cached per device (the reference rebuilds 26 Keras models per request,
        deepdream.py:401-423); a later call with a deeper `upto` appends the missing layers, so the
        411 MB fc1 matrix is only ever built when a dense layer is actually visualised."""
        import os
        if device is None:
            device = int(os.environ.get("DCV_DEVICE", os.environ.get("LOCAL_RANK", "0")))
        names = [l.name for l in self.layers]
        want = len(self.layers) if (upto is None or upto not in names) else names.index(upto) + 1
        cached = self._engines.get(device)
        if cached is None:
            cached = self._engines[device] = [_engine.Engine(device), 0, None]
        eng, loaded, version = cached
        cur_version = tuple(getattr(l, "_version", 0) for l in self.layers[:max(loaded, want)])
        if version is not None and version[:loaded] != cur_version[:loaded]:
            eng.clear()                     # weights changed since upload: start over
            loaded = 0
        for l in self.layers[loaded:want]:
            if l.kind == "conv":
                W, b = l.get_weights()
                if not l.use_bias:
                    b = np.zeros_like(b)
                if l.activation_name not in (None, "linear", "relu"):
                    raise NotImplementedError("conv activation %r" % (l.activation_name,))
                eng.add_conv(l.name, W, b, relu=(l.activation_name == "relu"))
            elif l.kind == "pool":
                eng.add_pool(l.name)
            elif l.kind == "flatten":
                eng.add_flatten(l.name)
            elif l.kind == "dense":
                W, b = l.get_weights()
                eng.add_dense(l.name, W, b if l.use_bias else np.zeros_like(b), l.activation_name)
            elif l.kind == "input":
                continue
            else:
                raise SystemExit("Cannot handle this type of layer")      # deepdream.py:418-421
        loaded = max(loaded, want)
        cached[1] = loaded
        cached[2] = tuple(getattr(l, "_version", 0) for l in self.layers[:loaded])
        return eng
