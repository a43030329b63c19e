"""Diagnostic: engine vs autograd vs stock DDP, per-step parameter / local-gradient differences.
Launch with torch.distributed.run --nproc-per-node N."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import pytorch_distributed_mnist_b200 as pdm
from pytorch_distributed_mnist_b200.ops.cnn_engine import CnnEngine, NAMES

rank, world = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1))
dev = torch.device('cuda', int(os.environ.get('LOCAL_RANK', 0)))
torch.cuda.set_device(dev)
dist.init_process_group('nccl', init_method='env://', device_id=dev)
torch.manual_seed(7)
a = pdm.DistributedDataParallel(pdm.CNN().to(dev), device_ids=[dev.index])
torch.manual_seed(7)
b = pdm.DistributedDataParallel(pdm.CNN().to(dev), device_ids=[dev.index])
torch.manual_seed(7)
sm = pdm.CNN().to(dev)
s = torch.nn.parallel.DistributedDataParallel(sm, device_ids=[dev.index])
oa, ob, os_ = pdm.FusedAdam(a.parameters(), lr=1e-3), pdm.FusedAdam(b.parameters(), lr=1e-3), torch.optim.Adam(s.parameters(), lr=1e-3)
for m in (a, b, s):
    m.eval()
eng = CnnEngine(b.module, b.flat, ob, ddp=b, max_batch=32)
if rank == 0:
    print('buckets', eng._ranges, 'offsets', list(b.flat.offsets), flush=True)
g = torch.Generator().manual_seed(70 + rank)
for it in range(int(os.environ.get('STEPS', 4))):
    x = torch.randn(32, 1, 28, 28, generator=g).to(dev)
    y = torch.randint(0, 10, (32,), generator=g).to(dev)
    for m, o in ((a, oa), (s, os_)):
        loss = torch.nn.functional.cross_entropy(m(x), y)
        o.zero_grad()
        loss.backward()
        o.step()
    eng.train_step(x, y, seed=1)
    torch.cuda.synchronize()
    if rank == 0:
        for i, n in enumerate(NAMES):
            pa, pb, ps = a.flat.view_of('params', i), b.flat.view_of('params', i), dict(sm.named_parameters())[n]
            ga, gb = a.flat.view_of('grads', i), b.flat.view_of('grads', i)
            print('step %d %-13s |eng-auto| p %.3e g %.3e (gmax %.3e)  |auto-stock| %.3e  |eng-stock| %.3e' % (
                it, n, (pa - pb).abs().max().item(), (ga - gb).abs().max().item(), ga.abs().max().item(),
                (pa - ps).abs().max().item(), (pb - ps).abs().max().item()), flush=True)
        if it == int(os.environ.get('STEPS', 4)) - 1:
            pb, ps = b.flat.view_of('params', 2), sm.conv2.weight
            d = (pb - ps).abs()
            idx = torch.nonzero(d > 2e-5)
            print('conv2.w outliers vs stock:', idx.tolist()[:16])
            for j in idx.tolist()[:6]:
                t = tuple(j)
                print('  ', t, 'eng %.6e auto %.6e stock %.6e  grad eng %.3e auto %.3e stock(avg) %.3e' % (
                    pb[t].item(), a.flat.view_of('params', 2)[t].item(), ps[t].item(),
                    b.flat.view_of('grads', 2)[t].item(), a.flat.view_of('grads', 2)[t].item(), sm.conv2.weight.grad[t].item()))
eng.close(); a.close(); b.close()
dist.destroy_process_group()
