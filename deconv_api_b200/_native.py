"""ctypes binding of libdcv.so (the C-ABI declared in include/dcv.h).

There is no Python or CPU implementation behind these calls: if the library is missing it is built
in-tree with nvcc, and if that is impossible the import fails loudly.
"""
import ctypes
import os
import re

from . import build as _build

_c = ctypes
_LIB = None

P_F32 = _c.POINTER(_c.c_float)
P_F64 = _c.POINTER(_c.c_double)
P_U8 = _c.POINTER(_c.c_uint8)
P_I32 = _c.POINTER(_c.c_int)

# name -> (restype, argtypes); must list every DCV_API symbol of include/dcv.h (tests check this)
SIGNATURES = {
    "dcv_version": (_c.c_int, []),
    "dcv_create": (_c.c_int, [_c.c_int, _c.POINTER(_c.c_void_p)]),
    "dcv_destroy": (None, [_c.c_void_p]),
    "dcv_last_error": (_c.c_char_p, [_c.c_void_p]),
    "dcv_device": (_c.c_int, [_c.c_void_p]),
    "dcv_sync": (_c.c_int, [_c.c_void_p]),
    "dcv_launch_count": (_c.c_longlong, [_c.c_void_p]),
    "dcv_stream": (_c.c_void_p, [_c.c_void_p]),
    "dcv_host_alloc": (_c.c_int, [_c.c_void_p, _c.c_size_t, _c.POINTER(_c.c_void_p)]),
    "dcv_host_free": (_c.c_int, [_c.c_void_p, _c.c_void_p]),
    "dcv_profile": (_c.c_int, [_c.c_void_p, _c.c_int]),
    "dcv_profile_read": (_c.c_int, [_c.c_void_p, _c.c_int, P_F64, _c.POINTER(_c.c_longlong), P_F64, P_F64]),
    "dcv_profile_launch": (_c.c_int, [_c.c_void_p, _c.c_int, P_I32, P_F64, P_F64, P_F64]),
    "dcv_net_clear": (_c.c_int, [_c.c_void_p]),
    "dcv_net_add_conv": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int]),
    "dcv_net_add_pool": (_c.c_int, [_c.c_void_p]),
    "dcv_net_add_flatten": (_c.c_int, [_c.c_void_p]),
    "dcv_net_add_dense": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int]),
    "dcv_net_num_layers": (_c.c_int, [_c.c_void_p]),
    "dcv_set_forward_mode": (_c.c_int, [_c.c_void_p, _c.c_int]),
    "dcv_set_fp32_mode": (_c.c_int, [_c.c_void_p, _c.c_int]),
    "dcv_forward": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int, _c.c_int]),
    "dcv_forward_dev": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int, _c.c_int]),
    "dcv_layer_shape": (_c.c_int, [_c.c_void_p, _c.c_int, P_I32]),
    "dcv_read_activation": (_c.c_int, [_c.c_void_p, _c.c_int, _c.c_void_p]),
    "dcv_read_switch": (_c.c_int, [_c.c_void_p, _c.c_int, _c.c_void_p]),
    "dcv_channel_sums": (_c.c_int, [_c.c_void_p, _c.c_int, _c.c_void_p]),
    "dcv_channel_sums_buf": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_size_t, _c.c_int, _c.c_void_p]),
    "dcv_project": (_c.c_int, [_c.c_void_p, _c.c_int, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int, _c.c_void_p]),
    "dcv_project_dev": (_c.c_int, [_c.c_void_p, _c.c_int, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int, _c.c_void_p]),
    "dcv_project_fmt": (_c.c_int, [_c.c_void_p, _c.c_int, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_void_p]),
    "dcv_project_fmt_dev": (_c.c_int, [_c.c_void_p, _c.c_int, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_void_p]),
    "dcv_request_mosaic": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_void_p,
                                      _c.c_void_p]),
    "dcv_deprocess": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_size_t, _c.c_void_p, _c.c_int]),
    "dcv_deprocess_dev": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_size_t, _c.c_void_p]),
    "dcv_deprocess_dev_noalloc": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_size_t, _c.c_void_p]),
    "dcv_maxpool_switch": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_void_p, _c.c_void_p]),
    "dcv_unpool_switch": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_void_p]),
    "dcv_conv3x3": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int,
                               _c.c_int, _c.c_int, _c.c_int, _c.c_void_p]),
    "dcv_deconv3x3": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int, _c.c_int,
                                 _c.c_int, _c.c_int, _c.c_int, _c.c_void_p]),
    "dcv_deconv_tail": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int, _c.c_void_p, _c.c_int, _c.c_int,
                                   _c.c_int, _c.c_void_p, _c.c_int, _c.c_int, _c.c_void_p]),
    "dcv_activation": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int, _c.c_void_p]),
    "dcv_dense": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int,
                             _c.c_int, _c.c_void_p]),
    "dcv_dense_transpose": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int, _c.c_void_p]),
    "dcv_dream": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int, _c.c_void_p, _c.c_void_p, _c.c_int,
                             _c.c_void_p, _c.c_int, _c.c_int, _c.c_float, _c.c_void_p, _c.c_int, _c.c_void_p, _c.c_void_p]),
    "dcv_dream_dev": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int, _c.c_void_p, _c.c_void_p, _c.c_int,
                                 _c.c_void_p, _c.c_int, _c.c_int, _c.c_float, _c.c_void_p, _c.c_int, _c.c_void_p]),
    "dcv_dream_gradient": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int, _c.c_void_p, _c.c_void_p,
                                      _c.c_int, _c.c_int, _c.c_void_p, _c.c_void_p]),
    "dcv_resize_bilinear": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_void_p]),
    "dcv_roll": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_void_p]),
    "dcv_maxpool_switch_dev": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_void_p, _c.c_void_p]),
    "dcv_unpool_switch_dev": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_void_p]),
    "dcv_resize_bilinear_dev": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_void_p]),
    "dcv_roll_dev": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_void_p]),
}

HEADER = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "dcv.h")


def header_symbols(path=HEADER):
    """Names of every DCV_API function declared in include/dcv.h."""
    with open(path) as f:
        return re.findall(r"^DCV_API\s+[\w\s\*]+?\b(dcv_\w+)\s*\(", f.read(), flags=re.M)


def lib_path():
    return _build.LIB


def load():
    """Load (building first if needed) the shared library and attach the signatures."""
    global _LIB
    if _LIB is not None:
        return _LIB
    override = os.environ.get("DCV_LIB_PATH")           # A/B measurements of differently compiled builds (profiles/)
    if override:
        lib = ctypes.CDLL(override)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _LIB = lib
        return lib
    # build() returns at once when the stored source hash matches, so a library left behind by an older
    # checkout of csrc/ is never loaded silently; without nvcc a stale or missing library is an error
    try:
        path = _build.build(force=bool(os.environ.get("DCV_REBUILD")))
    except Exception as e:
        raise ImportError("deconv_api_b200: %s is missing or older than csrc/ and cannot be rebuilt (%s); "
                          "there is no fallback" % (_build.LIB, e))
    try:
        lib = ctypes.CDLL(path)
    except OSError as e:
        raise ImportError("deconv_api_b200: cannot load the CUDA library %s (%s); there is no fallback" % (path, e))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)      # AttributeError = header/library mismatch: fail loudly
        fn.restype = res
        fn.argtypes = args
    _LIB = lib
    return lib


class DcvError(RuntimeError):
    def __init__(self, code, text):
        RuntimeError.__init__(self, "dcv error %d: %s" % (code, text))
        self.code = code


def check(lib, ctx, code):
    if code < 0:
        msg = lib.dcv_last_error(ctx)
        raise DcvError(code, msg.decode("utf-8", "replace") if msg else "")
    return code
