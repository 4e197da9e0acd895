"""Drop-in for reference ``app/deepdream.py``: the same classes and functions, same signatures,
argument meaning and error behaviour — with the arithmetic on the B200 through the C-ABI.

    reference symbol (app/deepdream.py)                 here
    ---------------------------------------------------------------------------------------------
    DInput / DConvolution2D / DPooling / DActivation    same names; .up/.down call one CUDA operator
    DDense / DFlatten                 (:13-366)         each on host ndarrays (per-call H2D/D2H)
    find_top_filters(output, top=8)   (:369-380)        per-channel fp64 sums on the GPU + the same
                                                        filter/sort rule on the <=512 sums
    visualize_all_layers(model, data, layer_name,       one forward tape kept on the device, then the
                         visualize_mode)  (:383-476)    projections of each layer batched in one call
    deprocess_image(x)                (:483-498)        two-pass reduction + fused map on the GPU

There is no CPU implementation behind any of them: without the CUDA library / a B200 they raise.
"""
import sys

import numpy as np

from . import engine as _engine
from . import layers as _layers

# Additive knobs (not in the reference): default arithmetic of the backward chain, and progress
# prints (the reference prints layer indices / shapes / top filters, deepdream.py:438,445-447).
DEFAULT_PRECISION = "fp32"
VERBOSE = False


def set_precision(p):
    global DEFAULT_PRECISION
    _engine.precision_id(p)
    DEFAULT_PRECISION = p


def _eng(device=None):
    return _engine.default_engine(device)


def _kind(layer):
    k = getattr(layer, "kind", None)
    if k:
        return k
    n = type(layer).__name__
    kinds = {"Conv2D": "conv", "Convolution2D": "conv", "MaxPooling2D": "pool", "Dense": "dense",
             "Flatten": "flatten", "Activation": "activation", "InputLayer": "input"}
    if n not in kinds:
        print("Cannot handle this type of layer")
        print(layer.get_config())
        sys.exit()                                     # deepdream.py:418-421
    return kinds[n]


def _act_name(layer):
    a = getattr(layer, "activation_name", None)
    if a is None:
        a = getattr(layer, "activation", None)
        a = getattr(a, "__name__", a)
    return a


def adopt(model):
    """Accept any Keras-shaped model (``.layers`` with name / get_weights / get_config …): returns a
    ``layers.Model`` description of it, cached on the object."""
    if isinstance(model, _layers.Model):
        return model
    cached = getattr(model, "_dcv_model", None)
    if cached is not None:
        return cached
    out = []
    for l in model.layers:
        k = _kind(l)
        if k == "input":
            out.append(_layers.InputLayer(tuple(l.input_shape[1:]), name=l.name))
        elif k == "conv":
            cfg = l.get_config()
            if tuple(cfg.get("strides", (1, 1))) != (1, 1):
                raise NotImplementedError("strided convolution")
            nl = _layers.Convolution2D(cfg["filters"], cfg["kernel_size"], activation=_act_name(l),
                                       padding=cfg.get("padding", "same"), use_bias=cfg.get("use_bias", True), name=l.name)
            nl._weights = [np.asarray(w, np.float32) for w in l.get_weights()]
            out.append(nl)
        elif k == "pool":
            out.append(_layers.MaxPooling2D(tuple(l.pool_size), name=l.name))
        elif k == "flatten":
            out.append(_layers.Flatten(name=l.name))
        elif k == "dense":
            cfg = l.get_config()
            out.append(_layers.Dense(cfg["units"], activation=_act_name(l), use_bias=cfg.get("use_bias", True),
                                     name=l.name, lazy=(lambda _l=l: _l.get_weights())))
        elif k == "activation":
            raise NotImplementedError("stand-alone Activation layers are not part of the VGG16 path")
    m = _layers.Model(out, name=getattr(model, "name", "model"))
    try:
        model._dcv_model = m
    except Exception:
        pass
    return m


# ------------------------------------------------------------------------------------------------
# Layer wrappers (reference deepdream.py:13-366).  Same constructor argument (a layer object), same
# .up(data)/.down(data) contract and the same public attributes (up_data, down_data, switch,
# up_func, down_func).  Each call is a single CUDA operator on host arrays.
class DInput(object):
    """deepdream.py:13-50 — identity in both directions."""

    def __init__(self, layer):
        self.layer = layer

    def up(self, data):
        self.up_data = data
        return self.up_data

    def down(self, data):
        self.down_data = data
        return self.down_data


class DConvolution2D(object):
    """deepdream.py:53-111.  up = act(conv(x, W) + b); down = act(conv(x, flipT(W))) with zero bias —
    the down layer is built from the forward layer's config, activation included (:69,86)."""

    def __init__(self, layer):
        self.layer = layer
        W, b = layer.get_weights()
        cfg = layer.get_config()
        self._W = np.ascontiguousarray(W, np.float32)
        self._b = np.ascontiguousarray(b, np.float32) if cfg.get("use_bias", True) else np.zeros(W.shape[3], np.float32)
        self._relu = (cfg.get("activation") == "relu")
        if cfg.get("activation") not in (None, "linear", "relu"):
            raise NotImplementedError("conv activation %r" % (cfg.get("activation"),))
        # the callables the reference keeps as Keras models (deepdream.py:76,89)
        self.up_func = lambda x: _eng().conv3x3(x, self._W, self._b, relu=self._relu)
        self.down_func = lambda x, precision=None: _eng().deconv3x3(x, self._W, relu=self._relu,
                                                                    precision=precision or DEFAULT_PRECISION)

    def up(self, data):
        self.up_data = self.up_func(data)
        return self.up_data

    def down(self, data):
        self.down_data = self.down_func(data)
        return self.down_data


class DPooling(object):
    """deepdream.py:114-209.  up: 2x2 max pool; ``switch`` is the reference's float64 one-hot mask
    (built from the kernel's 1-byte index), ``up_data`` is float64 like the reference's (:162,168)."""

    def __init__(self, layer):
        self.layer = layer
        self.poolsize = layer.pool_size
        if tuple(int(p) for p in self.poolsize) != (2, 2):
            raise NotImplementedError("only 2x2 pooling")

    def up(self, data):
        pooled, idx = _eng().maxpool_switch(data)
        self.switch_index = idx
        self.up_data = pooled.astype(np.float64)
        self.switch = switch_mask(idx, np.asarray(data).shape)
        self._mask_of_index = self.switch
        return self.up_data

    def down(self, data):
        sw = self.switch
        if sw is not getattr(self, "_mask_of_index", None):      # caller replaced .switch by hand
            self.switch_index = switch_index_from_mask(sw)
            self._mask_of_index = sw
        if (sw.shape[1] % 2) or (sw.shape[2] % 2):
            # the reference dies here with a numpy broadcast error
            raise ValueError("operands could not be broadcast together: unpool needs even spatial sizes, got %s" % (sw.shape,))
        self.down_data = _eng().unpool_switch(data, self.switch_index).astype(np.float64)
        return self.down_data


def switch_mask(idx, in_shape):
    """Expand the 1-byte switch index (B,H/2,W/2,C) into the reference's one-hot float64 mask of the
    pool input's shape (deepdream.py:162,184-187).  Pure re-encoding, no arithmetic."""
    B, H, W, C = in_shape
    oh, ow = H // 2, W // 2
    sw = np.zeros(in_shape, np.float64)
    for k in range(4):
        r, q = divmod(k, 2)
        sw[:, r:2 * oh:2, q:2 * ow:2, :] = (idx == k)
    return sw


def switch_index_from_mask(sw):
    B, H, W, C = sw.shape
    oh, ow = H // 2, W // 2
    idx = np.zeros((B, oh, ow, C), np.uint8)
    for k in range(1, 4):
        r, q = divmod(k, 2)
        idx[sw[:, r:2 * oh:2, q:2 * ow:2, :] != 0] = k
    return idx


class DActivation(object):
    """deepdream.py:212-261 — the layer's activation, applied identically in both directions."""

    def __init__(self, layer, linear=False):
        self.layer = layer
        self.linear = linear
        self.activation = layer.activation
        self._name = _act_name(layer)
        if self._name not in (None, "linear", "relu", "softmax"):
            raise NotImplementedError("activation %r" % (self._name,))
        self.up_func = self.down_func = self._apply

    def _apply(self, args):
        data = args[0] if isinstance(args, (list, tuple)) else args
        data = np.ascontiguousarray(data, np.float32)
        if self._name == "relu":
            return [_eng().relu(data)]
        if self._name == "softmax":
            return [_eng().softmax(data)]
        return [data]

    def up(self, data, learning_phase=0):
        self.up_data = self.up_func([data, learning_phase])[0]
        return self.up_data

    def down(self, data, learning_phase=0):
        self.down_data = self.down_func([data, learning_phase])[0]
        return self.down_data


class DDense(object):
    """deepdream.py:264-321.  up = act(x W + b); down = x W^T, zero bias, no activation (:289-297)."""

    def __init__(self, layer):
        self.layer = layer
        W, b = layer.get_weights()
        self._W = np.ascontiguousarray(W, np.float32)
        self._b = np.ascontiguousarray(b, np.float32)
        self._act = _act_name(layer)
        self.input_shape = layer.input_shape
        self.output_shape = layer.output_shape
        self.up_func = lambda x: _eng().dense(x, self._W, self._b, self._act)
        self.down_func = lambda x: _eng().dense_transpose(x, self._W)

    def up(self, data):
        self.up_data = self.up_func(data)
        return self.up_data

    def down(self, data):
        self.down_data = self.down_func(data)
        return self.down_data


class DFlatten(object):
    """deepdream.py:324-366 — reshape only."""

    def __init__(self, layer):
        self.layer = layer
        self.shape = layer.input_shape[1:]
        self.up_func = lambda args: [np.reshape(args[0], (np.asarray(args[0]).shape[0], -1))]

    def up(self, data, learning_phase=0):
        self.up_data = self.up_func([data, learning_phase])[0]
        return self.up_data

    def down(self, data):
        new_shape = [data.shape[0]] + list(self.shape)
        assert np.prod(self.shape) == np.prod(data.shape[1:])
        self.down_data = np.reshape(data, new_shape)
        return self.down_data


# ------------------------------------------------------------------------------------------------
def _top_from_sums(sums, top, dtype):
    """The selection rule of find_top_filters (deepdream.py:376-380) on the per-channel sums: keep
    sums > 0, stable sort descending, first `top`.  Returns [(index, sum)] like the reference."""
    filter_sum = [(i, dtype(s)) for i, s in enumerate(sums) if s > 0]
    filter_sum.sort(key=lambda x: x[1], reverse=True)
    return filter_sum[:top]


def find_top_filters(output, top=8):
    """deepdream.py:369-380.  The sums are accumulated in fp64 on the GPU (the reference sums in the
    array's own dtype: float32 after a conv, float64 after a pool)."""
    output = np.asarray(output)
    dtype = np.float64 if output.dtype == np.float64 else np.float32
    sums = _eng().channel_sums_array(output.reshape(-1, output.shape[-1]))
    return _top_from_sums(sums, top, dtype)


def visualize_all_layers(model, data, layer_name='predictions', visualize_mode='all',
                         precision=None, only_layers=None, top=8, device=None):
    '''
    function to visualize feature (reference deepdream.py:383-476)
    # Arguments
        model: model used to visualize data (``vgg16.VGG16(...)`` of this package, or any
               Keras-shaped model object)
        data: image(s) to visualize, (B,H,W,C) float32, already preprocessed
        layer_name: Name of layer to visualize; every model layer from this one down to the
                    first conv is projected, like the reference
        visualize_mode: 'all' or 'max' (deepdream.py:454-460)
      additive, not in the reference:
        precision: 'fp32' (parity, default) or 'bf16' (tensor cores) for the backward chain
        only_layers: restrict the projected layers to these names (main.py only uses one)
        top: number of feature maps per layer (reference: 8)
    # Returns
        dict: layer name -> list of (H,W,C) float32 projections (squeezed), strongest map first
    '''
    if visualize_mode not in ('all', 'max'):
        print('Illegal visualize mode')
        sys.exit()                                      # deepdream.py:458-460
    m = adopt(model)
    names = [l.name for l in m.layers]
    # the reference's stacking loop stops AFTER the layer called layer_name; an unknown name never
    # breaks, so every layer is stacked and visualised
    last = names.index(layer_name) if layer_name in names else len(names) - 1
    if last == 0:
        return dict()            # only the input layer was stacked; it is popped (deepdream.py:437)
    eng = m.bound_engine(device, upto=names[last])
    data = np.ascontiguousarray(data, dtype=np.float32)
    prec = precision or DEFAULT_PRECISION
    eng.forward(data, eng.index[names[last]] if names[last] in eng.index else len(eng.layer_names) - 1)

    deconv_dict = dict()
    # layers_to_visualize: model layers from the target down, input removed (deepdream.py:431-437)
    for li in range(last, 0, -1):
        name = names[li]
        if name not in eng.index:
            continue
        if only_layers is not None and name not in only_layers:
            continue
        idx = eng.index[name]
        kind = eng.kinds[idx]
        sums = eng.channel_sums(idx)
        # the reference sums float32 arrays at conv/dense layers and float64 ones at pool layers
        top_filters = _top_from_sums(sums, top, np.float64 if kind == "pool" else np.float32)
        if VERBOSE:
            print('output.shape :', eng.layer_shape(idx))
            print('deconv_layer:', name)
            print('top_filters:', top_filters)
        feats = [f for f, _ in top_filters]
        if feats:
            proj = eng.project(idx, feats, mode=visualize_mode, precision=prec)
            deconv_dict[name] = [proj[k].squeeze() for k in range(len(feats))]
        else:
            deconv_dict[name] = []
    return deconv_dict


def deprocess_image(x):
    """deepdream.py:483-498: centre, scale to std 0.1, shift to 0.5, clip, to uint8 — and, like the
    reference's in-place operators (:485-490), leaves the normalised values in ``x``."""
    if not (isinstance(x, np.ndarray) and x.dtype == np.float32 and x.flags.c_contiguous and x.flags.writeable):
        y = np.ascontiguousarray(x, dtype=np.float32)
        out = _eng().deprocess(y, mutate=True)
        try:
            x[...] = y
        except Exception:
            pass
        return out
    return _eng().deprocess(x, mutate=True)
