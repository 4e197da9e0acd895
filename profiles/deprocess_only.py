"""The de-process shape of profiles/stream_kernels.py alone (for ncu -k regex:deprocess)."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deconv_api_b200 import engine as E
eng = E.Engine(0)
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(0)
y = torch.randn((448 * 448 * 3 * 64,), device=dev, generator=g)
y8 = torch.empty(y.shape, device=dev, dtype=torch.uint8)
for _ in range(3):
    eng._ck(eng.lib.dcv_deprocess_dev(eng.ctx, ctypes.c_void_p(y.data_ptr()), y.numel(), ctypes.c_void_p(y8.data_ptr())))
eng.sync()
