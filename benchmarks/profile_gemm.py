#!/usr/bin/env python
"""Short single-GPU driver for ncu: the tcgen05 GEMM in its conv2-forward (width 8) and fc1 shapes."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from pytorch_distributed_mnist_b200.ops import gemm as G  # noqa: E402

dev = torch.device('cuda', 0)
B, c1, c2 = 128, 256, 512
a1 = torch.randn(B * 676, c1, device=dev)
wk = torch.randn(c2, 9 * c1, device=dev) * 0.05
z2 = torch.empty(B * 676, c2, device=dev)
for _ in range(3):
    G.gemm(a1, wk, out=z2, mode=G.MODE_CONV, img_w=26, cblocks=c1 // 32, M=B * 676, N=c2, K=9 * c1)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    G.gemm(a1, wk, out=z2, mode=G.MODE_CONV, img_w=26, cblocks=c1 // 32, M=B * 676, N=c2, K=9 * c1)
e1.record()
torch.cuda.synchronize()
us = e0.elapsed_time(e1) / 5 * 1e3
print('conv2 fwd implicit GEMM M={} N={} K={}: {:.1f} us, {:.1f} TFLOP/s'.format(B * 676, c2, 9 * c1, us, 2.0 * B * 676 * c2 * 9 * c1 / us / 1e6))
