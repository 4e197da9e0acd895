"""dcv_deprocess_dev_noalloc: the de-process entry point for callers that are not allowed to allocate GPU
memory (a server whose pool owns the device, a latency-critical request path, a sandbox with pre-declared
buffers only).  Its reduction scratch is a fixed array inside the library, so the call must give exactly what
dcv_deprocess_dev gives — on the first call (which also resolves the array's device address) and on a
following call that re-uses the array with stale contents from the first.

Named test_zz_* so that pytest collects it last: a run that stops at the first failure has then already
exercised everything else.  It passes alone in a fresh process and after the rest of the suite."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run(engine, fn, x_dev, n, out_dev):
    rc = fn(engine.ctx, ctypes.c_void_p(x_dev.data_ptr()), n, ctypes.c_void_p(out_dev.data_ptr()))
    assert rc == 0, engine.lib.dcv_last_error(engine.ctx)
    engine.sync()
    return out_dev.cpu().numpy()


def test_noalloc_deprocess_matches_regular_entry_point_twice(engine):
    import torch
    lib = engine.lib
    rng = np.random.default_rng(11)
    # two calls in a row: a large input first (fills many partials), then a small one (uses few of them,
    # the rest of the shared array still holds the first call's values)
    for n, mean, std in ((448 * 448 * 3, 0.7, 2.5), (37 * 29 * 3 + 1, -3.0, 0.4)):
        x = rng.normal(mean, std, (n,)).astype(np.float32)
        x_dev = torch.from_numpy(x).cuda()
        a = _run(engine, lib.dcv_deprocess_dev, x_dev, n, torch.zeros(n, dtype=torch.uint8, device="cuda"))
        b = _run(engine, lib.dcv_deprocess_dev_noalloc, x_dev, n, torch.full((n,), 77, dtype=torch.uint8, device="cuda"))
        assert np.array_equal(a, b)                      # same kernels, same partial order: identical bytes
        ref = np.clip(np.clip((x - x.mean()) / (x.std() + 1e-7) * 0.1 + 0.5, 0, 1) * 255, 0, 255).astype(np.uint8)
        d = np.abs(b.astype(np.int16) - ref.astype(np.int16))
        assert d.max() <= 1 and np.count_nonzero(d) <= max(1, n // 500)   # tolerance of the de-process tests
        assert torch.equal(x_dev.cpu(), torch.from_numpy(x))            # the input is not modified


def test_noalloc_call_leaves_free_memory_unchanged(engine):
    """The point of the entry point: once the first call on the device has resolved the array's address (and the
    runtime has loaded the kernels), a call must not change the amount of free device memory."""
    import torch
    lib = engine.lib
    n = 224 * 224 * 3 * 8
    x_dev = torch.randn(n, dtype=torch.float32, device="cuda")
    out_dev = torch.zeros(n, dtype=torch.uint8, device="cuda")
    _run(engine, lib.dcv_deprocess_dev_noalloc, x_dev, n, out_dev)          # warm-up: address lookup, module load
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        _run(engine, lib.dcv_deprocess_dev_noalloc, x_dev, n, out_dev)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0


def test_noalloc_two_contexts_share_the_array_in_stream_order():
    """Two contexts on one device use the same fixed array; called one after the other (each synchronised) both
    give the regular entry point's bytes."""
    import torch
    from deconv_api_b200 import engine as E
    a, b = E.Engine(0), E.Engine(0)
    try:
        rng = np.random.default_rng(5)
        for eng, n in ((a, 50001), (b, 1234), (a, 77)):
            x_dev = torch.from_numpy(rng.normal(1.0, 3.0, (n,)).astype(np.float32)).cuda()
            want = _run(eng, eng.lib.dcv_deprocess_dev, x_dev, n, torch.zeros(n, dtype=torch.uint8, device="cuda"))
            got = _run(eng, eng.lib.dcv_deprocess_dev_noalloc, x_dev, n, torch.zeros(n, dtype=torch.uint8, device="cuda"))
            assert np.array_equal(want, got)
    finally:
        a.close()
        b.close()


def test_noalloc_rejects_oversize_and_bad_pointers(engine):
    import torch
    lib = engine.lib
    x_dev = torch.zeros(64, dtype=torch.float32, device="cuda")
    out_dev = torch.zeros(64, dtype=torch.uint8, device="cuda")
    too_many = (1 << 28) + 1
    rc = lib.dcv_deprocess_dev_noalloc(engine.ctx, ctypes.c_void_p(x_dev.data_ptr()), too_many, ctypes.c_void_p(out_dev.data_ptr()))
    assert rc == -1
    msg = lib.dcv_last_error(engine.ctx)
    assert b"exceed" in msg and b"268435456" in msg
    assert lib.dcv_deprocess_dev_noalloc(engine.ctx, None, 64, ctypes.c_void_p(out_dev.data_ptr())) == -1
    assert lib.dcv_deprocess_dev_noalloc(engine.ctx, ctypes.c_void_p(x_dev.data_ptr()), 0, ctypes.c_void_p(out_dev.data_ptr())) == 0
    engine.sync()
