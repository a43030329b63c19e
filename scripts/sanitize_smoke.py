#!/usr/bin/env python
"""Smallest exercise of every native kernel family, for compute-sanitizer (one tool per run)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

import pytorch_distributed_mnist_b200 as pdm  # noqa: E402
from pytorch_distributed_mnist_b200.ops import gemm as G  # noqa: E402
from pytorch_distributed_mnist_b200.ops.fused_step import NetStepper  # noqa: E402
from pytorch_distributed_mnist_b200.ops.engine import NetEngine  # noqa: E402

dev = torch.device('cuda', 0)
torch.manual_seed(0)
net = pdm.Net().to(dev)
opt = pdm.FusedAdam(net.parameters(), lr=1e-3)
st = NetStepper(dev)
for B in (1, 24, 64):
    x = torch.randint(0, 256, (B, 784), dtype=torch.uint8, device=dev)
    y = torch.randint(0, 10, (B,), dtype=torch.uint8, device=dev)
    st.train_step(x, y, opt.flat.params, opt.flat.exp_avg, opt.flat.exp_avg_sq, opt.hyper(1))
    st.eval_step(x, y, opt.flat.params)
eng = NetEngine(opt.flat, opt, max_batch=64, slots=4, workers=2)
eng.set_dataset(torch.randint(0, 256, (300, 784), dtype=torch.uint8), torch.randint(0, 10, (300,), dtype=torch.uint8))
eng.train(torch.randperm(300), 64)
eng.place_dataset('device')
eng.train(torch.randperm(300), 64)
eng.evaluate(torch.arange(300), 64)
eng.close()
for a_mn, b_mn in ((False, False), (False, True), (True, True)):
    A, Bm = torch.randn(132, 64, device=dev), torch.randn(48, 64, device=dev)
    a = A.t().contiguous() if a_mn else A
    b = Bm.t().contiguous() if b_mn else Bm
    out = G.gemm(a, b, a_mn_major=a_mn, b_mn_major=b_mn)
    assert torch.allclose(out, A @ Bm.t(), rtol=2e-2, atol=2e-2)
model = pdm.CNN().to(dev).train()
copt = pdm.FusedAdam(model.parameters(), lr=1e-3)
xb = torch.randn(4, 1, 28, 28, device=dev)
yb = torch.randint(0, 10, (4,), device=dev)
loss = F.cross_entropy(model(xb), yb)
copt.zero_grad()
loss.backward()
copt.step()
torch.cuda.synchronize()
print('sanitize smoke ok, loss', float(loss))
