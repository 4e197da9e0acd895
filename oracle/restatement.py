"""TEST INFRASTRUCTURE ONLY — CPU restatement of the reference's deconvnet algorithm.

This is the oracle that travels to the GPU box (where /root/reference does not exist).  It is a
vectorised numpy / torch-CPU restatement of reference ``app/deepdream.py``; every function cites
the lines it follows.  It is PINNED two ways (tests/test_oracle.py):
  * against the reference's own file run unmodified on ``oracle.keras_shim`` (build container only);
  * against golden vectors produced by that run and committed under tests/golden/.
The conv / dense arithmetic itself lives in third-party Keras 2.3.1 + TensorFlow 1.13.0rc2
(reference app/requirements.txt:54,118), which are absent here; it is restated from the Keras
channels_last contract (HWIO kernel, cross-correlation, padding='same', stride 1, fp32).  The
reference has no tests or golden vectors of its own, so parity at that boundary is pinned only by
the committed fixtures.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product path never does.
"""
import numpy as np
import torch
import torch.nn.functional as F


# ----------------------------------------------------------------------------- layer ops
def _nchw(x):
    return torch.from_numpy(np.ascontiguousarray(np.asarray(x, np.float32))).permute(0, 3, 1, 2)


def _nhwc(t):
    return t.permute(0, 2, 3, 1).contiguous().numpy()


def conv_up(x, W, b, relu=True):
    """DConvolution2D.up (deepdream.py:91-100): the layer re-built from its own config
    (:69-76) = relu(conv3x3_same(x, W) + b).  x NHWC, W HWIO."""
    wt = torch.from_numpy(np.ascontiguousarray(W, np.float32)).permute(3, 2, 0, 1).contiguous()
    kh, kw = W.shape[0], W.shape[1]
    y = F.conv2d(_nchw(x), wt, torch.from_numpy(np.ascontiguousarray(b, np.float32)), padding=(kh // 2, kw // 2))
    if relu:
        y = F.relu(y)
    return _nhwc(y)


def conv_down(x, W, relu=True):
    """DConvolution2D.down (deepdream.py:102-111): W -> transpose(0,1,3,2) -> flip both spatial
    axes (:80-81), zero bias (:84), and — because the down layer is built from the forward layer's
    config (:69,86) — the same activation: relu(conv3x3_same(x, flipT(W)))."""
    Wd = np.transpose(W, (0, 1, 3, 2))[::-1, ::-1, :, :]
    return conv_up(x, np.ascontiguousarray(Wd), np.zeros((Wd.shape[3],), np.float32), relu=relu)


def maxpool_with_switch(x, poolsize=(2, 2)):
    """DPooling.__max_pooling_with_switch (deepdream.py:152-188), vectorised.

    Returns (pooled float64, switch float64 one-hot shaped like x, idx uint8) where idx is the
    row-major position (r*pw + c) of the one inside each patch.  Tie rule of :180-183: the first
    row whose row-max equals the patch max, then the first column inside that row = the first
    maximum in row-major order.  Output dims are floor-divided (:166-167); trailing rows/cols of an
    odd input are never visited and stay 0 in the switch.
    """
    x = np.asarray(x)
    ph, pw = int(poolsize[0]), int(poolsize[1])
    n, h, w, c = x.shape
    oh, ow = h // ph, w // pw
    patches = x[:, :oh * ph, :ow * pw, :].reshape(n, oh, ph, ow, pw, c).transpose(0, 1, 3, 5, 2, 4)
    flat = patches.reshape(n, oh, ow, c, ph * pw)
    idx = flat.argmax(axis=-1)                       # numpy argmax = first occurrence
    pooled = np.take_along_axis(flat, idx[..., None], axis=-1)[..., 0].astype(np.float64)
    onehot = np.zeros(flat.shape, np.float64)
    np.put_along_axis(onehot, idx[..., None], 1.0, axis=-1)
    sw = onehot.reshape(n, oh, ow, c, ph, pw).transpose(0, 1, 4, 2, 5, 3).reshape(n, oh * ph, ow * pw, c)
    switch = np.zeros(x.shape, np.float64)
    switch[:, :oh * ph, :ow * pw, :] = sw
    return pooled, switch, idx.astype(np.uint8)


def unpool_with_switch(x, switch):
    """DPooling.__max_unpooling_with_switch (deepdream.py:191-209): kron(x, ones(tile)) * switch."""
    x = np.asarray(x)
    th, tw = switch.shape[1] // x.shape[1], switch.shape[2] // x.shape[2]
    up = np.repeat(np.repeat(x, th, axis=1), tw, axis=2).astype(np.float64)
    return up * switch


def unpool_with_idx(x, idx, out_hw=None):
    """Same result as ``unpool_with_switch`` driven by the compact 2-bit index instead of the mask."""
    x = np.asarray(x)
    n, oh, ow, c = x.shape
    h, w = out_hw if out_hw else (2 * oh, 2 * ow)
    out = np.zeros((n, h, w, c), x.dtype)
    for k in range(4):
        r, q = divmod(k, 2)
        out[:, r:2 * oh:2, q:2 * ow:2, :] = np.where(idx == k, x, 0)
    return out


def find_top_filters(output, top=8):
    """find_top_filters (deepdream.py:369-380), same numpy calls so the sums are bit-identical:
    per-channel np.sum over the strided view, keep sums > 0, stable sort descending, first `top`."""
    filter_sum = []
    for f in range(output.shape[-1]):
        s = np.sum(output[:, f]) if output.ndim == 2 else np.sum(output[:, :, :, f])
        if s > 0:
            filter_sum.append((f, s))
    filter_sum.sort(key=lambda t: t[1], reverse=True)
    return filter_sum[:top]


def deprocess_image(x):
    """deprocess_image (deepdream.py:483-498), channels_last; mutates x like the original
    (the in-place -=, /=, *=, += of :485-490)."""
    x -= x.mean()
    x /= (x.std() + 1e-7)
    x *= 0.1
    x += 0.5
    x = np.clip(x, 0, 1)
    x *= 255
    return np.clip(x, 0, 255).astype("uint8")


def softmax(v):
    v = np.asarray(v, np.float32)
    e = np.exp(v - v.max(axis=-1, keepdims=True))
    return (e / e.sum(axis=-1, keepdims=True)).astype(np.float32)


# ----------------------------------------------------------------------------- driver
def _kind(layer):
    k = getattr(layer, "kind", None)
    if k:
        return k
    n = type(layer).__name__
    return {"Conv2D": "conv", "Convolution2D": "conv", "MaxPooling2D": "pool", "Dense": "dense",
            "Flatten": "flatten", "Activation": "activation", "InputLayer": "input"}[n]


class _Entry(object):
    __slots__ = ("name", "kind", "W", "b", "act", "up", "switch", "is_model_layer", "shape")

    def __init__(self, name, kind, is_model_layer, W=None, b=None, act=None):
        self.name, self.kind, self.is_model_layer = name, kind, is_model_layer
        self.W, self.b, self.act = W, b, act
        self.up = self.switch = self.shape = None


def _act_name(layer):
    a = getattr(layer, "activation", None)
    return getattr(a, "__name__", a)


def _apply_act(name, v):
    if name == "relu":
        return np.maximum(np.asarray(v, np.float32), 0)
    if name == "softmax":
        return softmax(v)
    return np.asarray(v, np.float32)


def build_entries(model, layer_name):
    """The wrapper stack of visualize_all_layers (deepdream.py:401-423): conv/dense layers get a
    second '<name>_activation' entry; the loop stops after the layer called `layer_name`."""
    entries = []
    for layer in model.layers:
        k = _kind(layer)
        if k in ("conv", "dense"):
            W, b = layer.get_weights()
            entries.append(_Entry(layer.name, k, True, W, b, _act_name(layer)))
            entries.append(_Entry(layer.name + "_activation", "activation", False, act=_act_name(layer)))
        elif k == "pool":
            entries.append(_Entry(layer.name, "pool", True))
        elif k == "activation":
            entries.append(_Entry(layer.name, "activation", True, act=_act_name(layer)))
        elif k in ("flatten", "input"):
            entries.append(_Entry(layer.name, k, True))
        else:
            raise SystemExit("Cannot handle this type of layer")        # deepdream.py:418-421
        if layer_name == layer.name:
            break
    return entries


def forward(entries, data):
    """Forward pass (deepdream.py:426-428).  Pool outputs are float64 like the reference's
    (:162,168); every Keras predict casts back to float32."""
    x = data
    for e in entries:
        if e.kind == "input":
            y = x
        elif e.kind == "conv":
            y = _apply_act(e.act, conv_up(x, e.W, e.b, relu=False))
        elif e.kind == "dense":
            y = _apply_act(e.act, (torch.from_numpy(np.ascontiguousarray(x, np.float32))
                                   @ torch.from_numpy(e.W) + torch.from_numpy(e.b)).numpy())
        elif e.kind == "activation":
            y = _apply_act(e.act, x)
        elif e.kind == "pool":
            y, e.switch, _ = maxpool_with_switch(x)
        elif e.kind == "flatten":
            e.shape = x.shape[1:]
            y = np.asarray(x).reshape(x.shape[0], -1)
        e.up = y
        x = y
    return entries


def backward(entries, i, seed):
    """One projection (deepdream.py:468-472): entry i's own down(), then every entry below it."""
    g = seed
    for j in range(i, -1, -1):
        e = entries[j]
        if e.kind == "conv":
            g = _apply_act(e.act, conv_down(g, e.W, relu=False))         # :102-111 (+ inherited act)
        elif e.kind == "dense":
            g = (torch.from_numpy(np.ascontiguousarray(g, np.float32))
                 @ torch.from_numpy(np.ascontiguousarray(e.W.T))).numpy()   # :289-297 linear, zero bias
        elif e.kind == "activation":
            g = _apply_act(e.act, g)                                      # :260
        elif e.kind == "pool":
            g = unpool_with_switch(g, e.switch)                           # :149
        elif e.kind == "flatten":
            g = np.reshape(g, [g.shape[0]] + list(e.shape))               # :363-365
        # input: identity (:40-50)
    return np.asarray(g).squeeze()


def visualize_all_layers(model, data, layer_name="predictions", visualize_mode="all", top=8,
                         only_layers=None):
    """visualize_all_layers (deepdream.py:383-476).  `only_layers` (not in the reference) restricts
    which layers are projected, to keep test time bounded; None = reference behaviour."""
    if visualize_mode not in ("all", "max"):
        raise SystemExit("Illegal visualize mode")                        # :458-460
    entries = forward(build_entries(model, layer_name), data)
    vis = [i for i, e in enumerate(entries) if e.is_model_layer]
    vis.reverse()
    vis.pop()                                                             # drop the input layer (:437)
    out = dict()
    for i in vis:
        e = entries[i]
        if only_layers is not None and e.name not in only_layers:
            continue
        output = e.up
        lst = []
        for f, _s in find_top_filters(output, top):
            assert output.ndim in (2, 4)
            fmap = output[:, f] if output.ndim == 2 else output[:, :, :, f]
            if visualize_mode == "max":
                fmap = fmap * (fmap == fmap.max())                        # :454-457, ties kept
            seed = np.zeros_like(output)
            if output.ndim == 2:
                seed[:, f] = fmap
            else:
                seed[:, :, :, f] = fmap
            lst.append(backward(entries, i, seed))
        out[e.name] = lst
    return out


# ----------------------------------------------------------------------------- strong CPU baseline
class FastTape(object):
    """Vectorised torch-CPU form of the same algorithm (F.conv2d / F.max_pool2d(return_indices) /
    F.conv_transpose2d / F.max_unpool2d), used as the timed CPU baseline and for large-size checks.
    tests/test_oracle.py proves it equal to the faithful restatement above (switches included)."""

    def __init__(self, model, layer_name, data):
        self.ops = []          # ("conv", W_oihw, b) / ("pool", indices, in_hw)
        x = _nchw(data)
        for layer in model.layers:
            k = _kind(layer)
            if k == "conv":
                W, b = layer.get_weights()
                wt = torch.from_numpy(np.ascontiguousarray(W, np.float32)).permute(3, 2, 0, 1).contiguous()
                x = F.conv2d(x, wt, torch.from_numpy(np.ascontiguousarray(b, np.float32)), padding=1)
                if _act_name(layer) == "relu":
                    x = F.relu(x)
                self.ops.append(("conv", wt, _act_name(layer) == "relu"))
            elif k == "pool":
                hw = x.shape[2:]
                x, idx = F.max_pool2d(x, 2, 2, return_indices=True)
                self.ops.append(("pool", idx, hw))
            elif k != "input":
                raise NotImplementedError(k)
            if layer.name == layer_name:
                break
        self.top = x

    def project(self, feats):
        """-> (n, H, W, C0) float32 projections of the top layer's feature maps `feats` (image 0)."""
        n = len(feats)
        g = torch.zeros((n,) + tuple(self.top.shape[1:]))
        for k, f in enumerate(feats):
            g[k, f] = self.top[0, f]
        for op in reversed(self.ops):
            if op[0] == "conv":
                g = F.conv_transpose2d(g, op[1], padding=1)
                if op[2]:
                    g = F.relu(g)
            else:
                idx = op[1].expand(n, -1, -1, -1)
                g = F.max_unpool2d(g, idx, 2, 2, output_size=list(op[2]))
        return _nhwc(g)
