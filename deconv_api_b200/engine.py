"""Engine: one dcv context (one GPU) holding a network and its forward tape.

Thin host-side wrapper over the C-ABI (include/dcv.h).  It owns no arithmetic: every method is one
or two library calls on numpy host buffers (or raw device pointers for the *_dev variants).
"""
import ctypes
import threading

import numpy as np

from . import _native

FP32, BF16 = 0, 1
VIS = {"all": 0, "max": 1}
_PREC = {"fp32": FP32, "bf16": BF16, FP32: FP32, BF16: BF16}
ACT = {"linear": 0, None: 0, "relu": 1, "softmax": 2}
# result formats of project(): name -> (dcv_out_format id, numpy dtype of the returned array)
OUT_FORMATS = {"f32": (0, np.float32), "bf16": (1, np.uint16), "u8": (2, np.uint8)}


def bf16_bits_to_f32(a):
    """uint16 bf16 bit patterns (project(..., out_format='bf16')) -> float32 values."""
    return (np.asarray(a, np.uint16).astype(np.uint32) << 16).view(np.float32)

_default_lock = threading.Lock()
_default_engines = {}


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def precision_id(p):
    try:
        return _PREC[p]
    except KeyError:
        raise ValueError("precision must be 'fp32' or 'bf16', got %r" % (p,))


class Engine(object):
    def __init__(self, device=0):
        self.lib = _native.load()
        h = ctypes.c_void_p()
        rc = self.lib.dcv_create(int(device), ctypes.byref(h))
        if rc < 0:
            msg = self.lib.dcv_last_error(None)
            raise _native.DcvError(rc, "dcv_create(device=%d): %s (no CPU fallback exists)" % (device, msg.decode() if msg else ""))
        self.ctx = h
        self.device = int(device)
        self.layer_names = []       # net index -> model layer name
        self.index = {}             # model layer name -> net index
        self.kinds = []
        self._net_owner = None

    # -- life cycle
    def close(self):
        if getattr(self, "ctx", None):
            for addr in list(getattr(self, "_pinned", ())):
                self._pinned.discard(addr)
                self.lib.dcv_host_free(self.ctx, ctypes.c_void_p(addr))
            self.lib.dcv_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        return _native.check(self.lib, self.ctx, rc)

    def sync(self):
        self._ck(self.lib.dcv_sync(self.ctx))

    @property
    def launch_count(self):
        return int(self.lib.dcv_launch_count(self.ctx))

    @property
    def stream(self):
        return self.lib.dcv_stream(self.ctx)

    # -- page-locked host buffers
    def pinned_empty(self, shape, dtype):
        """numpy array in page-locked host memory (full-rate, overlappable device-to-host copies).  The memory is
        released when the array (and every view of it) has been garbage-collected, or at close()."""
        import weakref
        dtype = np.dtype(dtype)
        nbytes = int(np.prod(shape)) * dtype.itemsize
        p = ctypes.c_void_p()
        self._ck(self.lib.dcv_host_alloc(self.ctx, nbytes, ctypes.byref(p)))
        buf = (ctypes.c_uint8 * max(nbytes, 1)).from_address(p.value)
        arr = np.frombuffer(buf, dtype=np.uint8, count=nbytes).view(dtype).reshape(shape)
        lib, ctx, addr = self.lib, self.ctx, p.value
        live = self._pinned = getattr(self, "_pinned", set())
        live.add(addr)

        def release(_lib=lib, _ctx=ctx, _addr=addr, _live=live):
            if _addr in _live:
                _live.discard(_addr)
                _lib.dcv_host_free(_ctx, ctypes.c_void_p(_addr))
        weakref.finalize(buf, release)
        return arr

    # -- per-kernel-class timing
    KERNEL_CLASSES = ("conv_fp32", "conv_tc", "pool", "unpool", "rank1", "other", "conv_fwd")

    def profile(self, enable=True):
        self._ck(self.lib.dcv_profile(self.ctx, 1 if enable else 0))

    def profile_read(self):
        """-> {class: dict(ms, launches, flops, bytes)} for the launches since profile(True)."""
        out = {}
        for i, name in enumerate(self.KERNEL_CLASSES):
            ms, fl, by = ctypes.c_double(), ctypes.c_double(), ctypes.c_double()
            n = ctypes.c_longlong()
            self._ck(self.lib.dcv_profile_read(self.ctx, i, ctypes.byref(ms), ctypes.byref(n), ctypes.byref(fl), ctypes.byref(by)))
            out[name] = dict(ms=ms.value, launches=n.value, flops=fl.value, bytes=by.value)
        return out

    def profile_launches(self):
        """-> [(class, ms, flops, bytes)] per recorded launch, in launch order."""
        out, i = [], 0
        while True:
            cat = ctypes.c_int()
            ms, fl, by = ctypes.c_double(), ctypes.c_double(), ctypes.c_double()
            if self.lib.dcv_profile_launch(self.ctx, i, ctypes.byref(cat), ctypes.byref(ms), ctypes.byref(fl), ctypes.byref(by)) < 0:
                return out
            out.append((self.KERNEL_CLASSES[cat.value], ms.value, fl.value, by.value))
            i += 1

    # -- network
    def clear(self):
        self._ck(self.lib.dcv_net_clear(self.ctx))
        self.layer_names, self.index, self.kinds = [], {}, []
        self._net_owner = None
        self._in_channels = None

    def add_conv(self, name, W, b, relu=True):
        W = _f32(W)
        assert W.ndim == 4 and W.shape[0] == 3 and W.shape[1] == 3, "only 3x3 kernels (HWIO) are supported"
        cin, cout = W.shape[2], W.shape[3]
        b = _f32(b) if b is not None else np.zeros((cout,), np.float32)
        idx = self._ck(self.lib.dcv_net_add_conv(self.ctx, _ptr(W), _ptr(b), cin, cout, 1 if relu else 0))
        if not any(k == "conv" for k in self.kinds):
            self._in_channels = cin          # the C-ABI takes the input channel count from the first conv
        self._register(name, idx, "conv")
        return idx

    def add_pool(self, name):
        idx = self._ck(self.lib.dcv_net_add_pool(self.ctx))
        self._register(name, idx, "pool")
        return idx

    def add_flatten(self, name):
        idx = self._ck(self.lib.dcv_net_add_flatten(self.ctx))
        self._register(name, idx, "flatten")
        return idx

    def add_dense(self, name, W, b, activation=None):
        W = _f32(W)
        b = _f32(b) if b is not None else np.zeros((W.shape[1],), np.float32)
        idx = self._ck(self.lib.dcv_net_add_dense(self.ctx, _ptr(W), _ptr(b), W.shape[0], W.shape[1], ACT[activation]))
        self._register(name, idx, "dense")
        return idx

    def _register(self, name, idx, kind):
        assert idx == len(self.layer_names)
        self.layer_names.append(name)
        self.kinds.append(kind)
        self.index[name] = idx

    # -- forward tape
    def set_forward_mode(self, tensor_cores):
        """True (default): tcgen05 split-operand forward, fp32-accurate (~1e-6 per layer); False: FFMA fp32 forward."""
        self._ck(self.lib.dcv_set_forward_mode(self.ctx, 1 if tensor_cores else 0))

    def set_fp32_mode(self, tensor_cores):
        """Backward convs of the fp32 chain: True (default) split-operand tcgen05, False CUDA-core FFMA."""
        self._ck(self.lib.dcv_set_fp32_mode(self.ctx, 1 if tensor_cores else 0))

    def forward(self, data, upto):
        data = _f32(data)
        assert data.ndim == 4, "data must be (B,H,W,C)"
        want = getattr(self, "_in_channels", None)
        if want is not None and data.shape[3] != want:
            raise _native.DcvError(-1, "forward: the network expects %d input channels, data has %d" % (want, data.shape[3]))
        B, H, W, _ = data.shape
        self._ck(self.lib.dcv_forward(self.ctx, _ptr(data), B, H, W, int(upto)))
        self.input_shape = data.shape

    def forward_dev(self, dev_ptr, B, H, W, upto):
        self._ck(self.lib.dcv_forward_dev(self.ctx, ctypes.c_void_p(dev_ptr), B, H, W, int(upto)))

    def layer_shape(self, idx):
        dims = (ctypes.c_int * 4)()
        self._ck(self.lib.dcv_layer_shape(self.ctx, int(idx), dims))
        return tuple(dims)

    def activation(self, idx):
        shp = self.layer_shape(idx)
        out = np.empty(shp, np.float32)
        self._ck(self.lib.dcv_read_activation(self.ctx, int(idx), _ptr(out)))
        if self.kinds[idx] in ("dense", "flatten"):
            out = out.reshape(shp[0], shp[3])
        return out

    def switch_index(self, idx):
        out = np.empty(self.layer_shape(idx), np.uint8)
        self._ck(self.lib.dcv_read_switch(self.ctx, int(idx), _ptr(out)))
        return out

    def channel_sums(self, idx):
        out = np.empty((self.layer_shape(idx)[3],), np.float64)
        self._ck(self.lib.dcv_channel_sums(self.ctx, int(idx), _ptr(out)))
        return out

    def channel_sums_array(self, x):
        x = _f32(x)
        C = x.shape[-1]
        out = np.empty((C,), np.float64)
        self._ck(self.lib.dcv_channel_sums_buf(self.ctx, _ptr(x), x.size // C, C, _ptr(out)))
        return out

    # -- projection
    def project(self, idx, feats, mode="all", precision="fp32", out=None, out_format="f32"):
        """-> (n, B, H0, W0, C0) projections.  out_format 'f32' (what the reference returns), 'bf16' (uint16 bit
        patterns, see bf16_bits_to_f32) or 'u8' (every projection passed through deprocess_image on the GPU).
        `out` may be a preallocated (e.g. pinned) array of the matching dtype."""
        feats = np.ascontiguousarray(feats, dtype=np.int32)
        n = int(feats.size)
        fmt, dtype = OUT_FORMATS[out_format]
        shape = (n,) + tuple(self.input_shape)
        if out is None:
            out = np.empty(shape, dtype)
        assert out.dtype == dtype and out.size == int(np.prod(shape)) and out.flags.c_contiguous
        if fmt == 0:
            self._ck(self.lib.dcv_project(self.ctx, int(idx), _ptr(feats), n, VIS[mode], precision_id(precision), _ptr(out)))
        else:
            self._ck(self.lib.dcv_project_fmt(self.ctx, int(idx), _ptr(feats), n, VIS[mode], precision_id(precision), fmt,
                                              _ptr(out)))
        return out.reshape(shape)

    def project_dev(self, idx, feats, out_dev_ptr, mode="all", precision="fp32", out_format="f32"):
        feats = np.ascontiguousarray(feats, dtype=np.int32)
        fmt = OUT_FORMATS[out_format][0]
        if fmt == 0:
            self._ck(self.lib.dcv_project_dev(self.ctx, int(idx), _ptr(feats), int(feats.size), VIS[mode],
                                              precision_id(precision), ctypes.c_void_p(out_dev_ptr)))
        else:
            self._ck(self.lib.dcv_project_fmt_dev(self.ctx, int(idx), _ptr(feats), int(feats.size), VIS[mode],
                                                  precision_id(precision), fmt, ctypes.c_void_p(out_dev_ptr)))

    def request_mosaic(self, img_bgr_u8, idx, mode="all", precision="fp32"):
        """The request tail of reference main.py:60-72 in one call: (H,W,3) uint8 image as cv2 decoded it ->
        ((2H,2W,3) uint8 mosaic of the four strongest maps of layer `idx`, their feature indices).  Raises IndexError
        like main.py:67-68 when fewer than four maps have a positive sum."""
        img = np.ascontiguousarray(img_bgr_u8, dtype=np.uint8)
        assert img.ndim == 3 and img.shape[2] == 3, "image must be (H,W,3) uint8"
        H, W, _ = img.shape
        out = np.empty((2 * H, 2 * W, 3), np.uint8)
        feats = np.empty((4,), np.int32)
        rc = self.lib.dcv_request_mosaic(self.ctx, _ptr(img), H, W, int(idx), VIS[mode], precision_id(precision), _ptr(out),
                                         _ptr(feats))
        if rc == -3:
            msg = (self.lib.dcv_last_error(self.ctx) or b"").decode("utf-8", "replace")
            if msg.startswith("list index out of range"):
                raise IndexError(msg)
        self._ck(rc)
        self.input_shape = (1, H, W, 3)
        return out, feats

    # -- image tail
    def deprocess(self, x, mutate=True):
        """x float32 C-contiguous; returns uint8 of the same shape; mutates x like the reference."""
        assert x.dtype == np.float32 and x.flags.c_contiguous
        out = np.empty(x.shape, np.uint8)
        self._ck(self.lib.dcv_deprocess(self.ctx, _ptr(x), x.size, _ptr(out), 1 if mutate else 0))
        return out

    # -- standalone operators
    def maxpool_switch(self, x):
        x = _f32(x)
        B, H, W, C = x.shape
        pooled = np.empty((B, H // 2, W // 2, C), np.float32)
        sw = np.empty((B, H // 2, W // 2, C), np.uint8)
        self._ck(self.lib.dcv_maxpool_switch(self.ctx, _ptr(x), B, H, W, C, _ptr(pooled), _ptr(sw)))
        return pooled, sw

    def unpool_switch(self, x, sw):
        x = _f32(x)
        sw = np.ascontiguousarray(sw, dtype=np.uint8)
        B, h, w, C = x.shape
        assert sw.shape == x.shape
        out = np.empty((B, 2 * h, 2 * w, C), np.float32)
        self._ck(self.lib.dcv_unpool_switch(self.ctx, _ptr(x), _ptr(sw), B, 2 * h, 2 * w, C, _ptr(out)))
        return out

    def conv3x3(self, x, W, b=None, relu=True):
        x, W = _f32(x), _f32(W)
        B, H, Wd, cin = x.shape
        cout = W.shape[3]
        assert W.shape == (3, 3, cin, cout)
        bb = _f32(b) if b is not None else None
        out = np.empty((B, H, Wd, cout), np.float32)
        self._ck(self.lib.dcv_conv3x3(self.ctx, _ptr(x), _ptr(W), _ptr(bb) if bb is not None else None,
                                      B, H, Wd, cin, cout, 1 if relu else 0, _ptr(out)))
        return out

    def deconv3x3(self, x, W, relu=True, precision="fp32"):
        """W is the FORWARD kernel (3,3,cin,cout); x has cout channels; result has cin channels."""
        x, W = _f32(x), _f32(W)
        B, H, Wd, cout = x.shape
        cin = W.shape[2]
        assert W.shape == (3, 3, cin, cout)
        out = np.empty((B, H, Wd, cin), np.float32)
        self._ck(self.lib.dcv_deconv3x3(self.ctx, _ptr(x), _ptr(W), B, H, Wd, cin, cout, 1 if relu else 0,
                                        precision_id(precision), _ptr(out)))
        return out


    def deconv_tail(self, x, W2, W1=None, sw=None, relu2=True, relu1=True):
        """[unpool by `sw`] -> backward conv of forward kernel W2 (3,3,cmid,c2) -> [backward conv of W1 (3,3,c0,cmid)],
        the fused bf16 tail kernel.  x: (B,H,W,c2), or the pooled (B,H/2,W/2,c2) when sw is given."""
        x, W2 = _f32(x), _f32(W2)
        B, ih, iw, c2 = x.shape
        H, W = (2 * ih, 2 * iw) if sw is not None else (ih, iw)
        cmid = W2.shape[2]
        assert W2.shape == (3, 3, cmid, c2)
        c0 = 0
        if W1 is not None:
            W1 = _f32(W1)
            c0 = W1.shape[2]
            assert W1.shape == (3, 3, c0, cmid)
        if sw is not None:
            sw = np.ascontiguousarray(sw, np.uint8)
            assert sw.shape == x.shape
        out = np.empty((B, H, W, c0 if W1 is not None else cmid), np.float32)
        self._ck(self.lib.dcv_deconv_tail(self.ctx, _ptr(x), _ptr(sw) if sw is not None else None, B, H, W, _ptr(W2), cmid, c2,
                                          1 if relu2 else 0, _ptr(W1) if W1 is not None else None, c0, 1 if relu1 else 0, _ptr(out)))
        return out

    def _activation(self, x, act):
        x = _f32(x)
        n = x.shape[-1]
        out = np.empty(x.shape, np.float32)
        self._ck(self.lib.dcv_activation(self.ctx, _ptr(x), x.size // n, n, ACT[act], _ptr(out)))
        return out

    def relu(self, x):
        return self._activation(x, "relu")

    def softmax(self, x):
        return self._activation(x, "softmax")

    def dense(self, x, W, b=None, activation=None):
        x, W = _f32(x), _f32(W)
        B, nin = x.shape
        nout = W.shape[1]
        assert W.shape[0] == nin
        bb = _f32(b) if b is not None else None
        out = np.empty((B, nout), np.float32)
        self._ck(self.lib.dcv_dense(self.ctx, _ptr(x), _ptr(W), _ptr(bb) if bb is not None else None, B, nin, nout,
                                    ACT[activation], _ptr(out)))
        return out

    def dense_transpose(self, x, W):
        x, W = _f32(x), _f32(W)
        B, nout = x.shape
        nin = W.shape[0]
        assert W.shape[1] == nout
        out = np.empty((B, nin), np.float32)
        self._ck(self.lib.dcv_dense_transpose(self.ctx, _ptr(x), _ptr(W), B, nin, nout, _ptr(out)))
        return out


    # -- DeepDream (derived spec)
    def resize_bilinear(self, x, out_hw):
        x = _f32(x)
        B, H, W, C = x.shape
        out = np.empty((B, int(out_hw[0]), int(out_hw[1]), C), np.float32)
        self._ck(self.lib.dcv_resize_bilinear(self.ctx, _ptr(x), B, H, W, C, int(out_hw[0]), int(out_hw[1]), _ptr(out)))
        return out

    def roll(self, x, oy, ox):
        x = _f32(x)
        B, H, W, C = x.shape
        out = np.empty_like(x)
        self._ck(self.lib.dcv_roll(self.ctx, _ptr(x), B, H, W, C, int(oy), int(ox), _ptr(out)))
        return out

    # -- device-resident forms (raw device pointers, e.g. torch.Tensor.data_ptr(); asynchronous on the context stream)
    def maxpool_switch_dev(self, x_ptr, shape, pooled_ptr, sw_ptr):
        B, H, W, C = (int(v) for v in shape)
        self._ck(self.lib.dcv_maxpool_switch_dev(self.ctx, ctypes.c_void_p(x_ptr), B, H, W, C, ctypes.c_void_p(pooled_ptr),
                                                 ctypes.c_void_p(sw_ptr)))

    def unpool_switch_dev(self, x_ptr, sw_ptr, out_shape, out_ptr):
        B, H, W, C = (int(v) for v in out_shape)
        self._ck(self.lib.dcv_unpool_switch_dev(self.ctx, ctypes.c_void_p(x_ptr), ctypes.c_void_p(sw_ptr), B, H, W, C,
                                                ctypes.c_void_p(out_ptr)))

    def resize_bilinear_dev(self, x_ptr, shape, out_hw, out_ptr):
        B, H, W, C = (int(v) for v in shape)
        self._ck(self.lib.dcv_resize_bilinear_dev(self.ctx, ctypes.c_void_p(x_ptr), B, H, W, C, int(out_hw[0]), int(out_hw[1]),
                                                  ctypes.c_void_p(out_ptr)))

    def roll_dev(self, x_ptr, shape, oy, ox, out_ptr):
        B, H, W, C = (int(v) for v in shape)
        self._ck(self.lib.dcv_roll_dev(self.ctx, ctypes.c_void_p(x_ptr), B, H, W, C, int(oy), int(ox), ctypes.c_void_p(out_ptr)))

    def dream_gradient(self, img, loss_layers, coeffs, precision="fp32"):
        """-> (dloss/dimg (B,H,W,C) float32, loss float)"""
        img = _f32(img)
        B, H, W, _ = img.shape
        ll = np.ascontiguousarray(loss_layers, np.int32)
        cf = np.ascontiguousarray(coeffs, np.float32)
        g = np.empty_like(img)
        loss = ctypes.c_double()
        self._ck(self.lib.dcv_dream_gradient(self.ctx, _ptr(img), B, H, W, _ptr(ll), _ptr(cf), int(ll.size), precision_id(precision),
                                             _ptr(g), ctypes.byref(loss)))
        return g, loss.value

    def dream(self, img, loss_layers, coeffs, shapes, iterations, step, jitter, want_losses=False, precision="fp32"):
        img = _f32(img)
        B, H, W, _ = img.shape
        ll = np.ascontiguousarray(loss_layers, np.int32)
        cf = np.ascontiguousarray(coeffs, np.float32)
        sh = np.ascontiguousarray(shapes, np.int32).reshape(-1, 2)
        jt = np.ascontiguousarray(jitter, np.int32).reshape(-1, 2)
        assert jt.shape[0] == sh.shape[0] * iterations, "need one (oy, ox) pair per iteration"
        out = np.empty_like(img)
        losses = np.zeros((sh.shape[0] * iterations,), np.float64) if want_losses else None
        self._ck(self.lib.dcv_dream(self.ctx, _ptr(img), B, H, W, _ptr(ll), _ptr(cf), int(ll.size), _ptr(sh), int(sh.shape[0]),
                                    int(iterations), ctypes.c_float(step), _ptr(jt), precision_id(precision), _ptr(out),
                                    _ptr(losses) if losses is not None else None))
        return (out, losses) if want_losses else out

    def dream_dev(self, img_dev_ptr, B, H, W, loss_layers, coeffs, shapes, iterations, step, jitter, out_dev_ptr, precision="fp32"):
        ll = np.ascontiguousarray(loss_layers, np.int32)
        cf = np.ascontiguousarray(coeffs, np.float32)
        sh = np.ascontiguousarray(shapes, np.int32).reshape(-1, 2)
        jt = np.ascontiguousarray(jitter, np.int32).reshape(-1, 2)
        self._ck(self.lib.dcv_dream_dev(self.ctx, ctypes.c_void_p(img_dev_ptr), B, H, W, _ptr(ll), _ptr(cf), int(ll.size), _ptr(sh),
                                        int(sh.shape[0]), int(iterations), ctypes.c_float(step), _ptr(jt), precision_id(precision),
                                        ctypes.c_void_p(out_dev_ptr)))


def default_engine(device=None):
    """Process-wide engine for `device` (default: LOCAL_RANK or 0), created on first use."""
    import os
    if device is None:
        device = int(os.environ.get("DCV_DEVICE", os.environ.get("LOCAL_RANK", "0")))
    with _default_lock:
        eng = _default_engines.get(device)
        if eng is None:
            eng = _default_engines[device] = Engine(device)
        return eng
