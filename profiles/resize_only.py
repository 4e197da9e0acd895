"""The two resize shapes of profiles/stream_kernels.py alone (for ncu):
    ncu --set full --clock-control none -k regex:resize -c 4 -o gpurun_out/resize python profiles/resize_only.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deconv_api_b200 import engine as E
eng = E.Engine(0)
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(0)
img = torch.randn((8, 1462, 1462, 3), device=dev, generator=g)
big = torch.empty((8, 2048, 2048, 3), device=dev)
small = torch.empty((8, 1462, 1462, 3), device=dev)
for _ in range(2):
    eng.resize_bilinear_dev(img.data_ptr(), img.shape, (2048, 2048), big.data_ptr())
    eng.resize_bilinear_dev(big.data_ptr(), big.shape, (1462, 1462), small.data_ptr())
eng.sync()
