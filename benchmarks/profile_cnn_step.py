#!/usr/bin/env python
"""Short single-GPU driver for ncu: a few steps of the native CNN step executor (width 1, batch 512) --
the FMA kernels (conv1, pool, fc2) and the fused all-reduce + optimizer bucket kernel."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import pytorch_distributed_mnist_b200 as pdm  # noqa: E402
from pytorch_distributed_mnist_b200.ops.cnn_engine import CnnEngine  # noqa: E402

dev = torch.device('cuda', 0)
torch.manual_seed(0)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
model = pdm.CNN().to(dev)
ddp = pdm.DistributedDataParallel(model, device_ids=[0])
opt = pdm.FusedAdam(ddp.parameters(), lr=1e-3)
eng = CnnEngine(model, ddp.flat, opt, ddp=ddp, max_batch=B)
ddp.train()
x = torch.randint(0, 256, (B, 784), dtype=torch.uint8, device=dev)
y = torch.randint(0, 10, (B,), device=dev)
for i in range(steps):
    loss = eng.train_step(x, y, seed=i)
torch.cuda.synchronize()
print('cnn step ok, loss {:.4f}'.format(float(loss)))
eng.close()
ddp.close()
