#!/usr/bin/env python
"""CNN extension (BASELINE.json configs 2-4): training images/s, ours vs the stock PyTorch stack.

    python benchmarks/bench_cnn.py [--width 1] [--batch 512] [--steps 50] [--warmup 10]
    python -m torch.distributed.run --nproc-per-node N ... benchmarks/bench_cnn.py ...

engine: the same kernels driven by the native step executor (csrc/cnn_step.cu, one C call per step;
        the path Trainer.train() takes for --arch cnn).
ours  : models.CNN on the hand-written kernels (tcgen05 implicit-GEMM conv, fused pool/dropout,
        tcgen05 fc1, fused fc2), our DistributedDataParallel (peer-memory bucketed all-reduce fused
        with Adam), fused cross-entropy.
stock : the same nn.Module on cuDNN/cuBLAS (TF32 allowed for conv AND matmul, i.e. the same
        precision class as ours), torch DistributedDataParallel + torch.optim.Adam.
Both: one process per GPU, synthetic uint8->fp32 inputs resident on the device, CUDA-event timing,
max over ranks.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import torch.nn.functional as F  # noqa: E402


def run(kind, a, rank, local, world, dev):
    import pytorch_distributed_mnist_b200 as pdm
    from pytorch_distributed_mnist_b200.ops.loss import cross_entropy
    torch.manual_seed(0)
    model = pdm.CNN(width=a.width).to(dev)
    eng = None
    if kind in ('ours', 'engine'):
        ddp = pdm.DistributedDataParallel(model, device_ids=[local])
        opt = pdm.FusedAdam(ddp.parameters(), lr=1e-3)
        loss_fn = cross_entropy
        if kind == 'engine':
            # native step executor (csrc/cnn_step.cu): what Trainer.train() drives for --arch cnn
            from pytorch_distributed_mnist_b200.ops.cnn_engine import CnnEngine
            eng = CnnEngine(model, ddp.flat, opt, ddp=ddp, max_batch=a.batch)
    else:
        torch.backends.cuda.matmul.allow_tf32 = True
        torch.backends.cudnn.allow_tf32 = True
        torch.backends.cudnn.benchmark = True
        # force the stock ops even on CUDA
        from pytorch_distributed_mnist_b200.ops import cnn_ops
        cnn_ops.available = lambda: False
        ddp = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local])
        opt = torch.optim.Adam(ddp.parameters(), lr=1e-3)
        loss_fn = F.cross_entropy
    ddp.train()
    g = torch.Generator(device=dev).manual_seed(1 + rank)
    nb = 8
    xs = [torch.randn(a.batch, 1, 28, 28, device=dev, generator=g) for _ in range(nb)]
    ys = [torch.randint(0, 10, (a.batch,), device=dev, generator=g) for _ in range(nb)]

    def step(i):
        if eng is not None:
            return eng.train_step(xs[i % nb], ys[i % nb], seed=i)
        out = ddp(xs[i % nb])
        loss = loss_fn(out, ys[i % nb])
        opt.zero_grad()
        loss.backward()
        opt.step()
        return loss

    for i in range(a.warmup):
        step(i)
    torch.cuda.synchronize(dev)
    dist.barrier()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(a.steps):
        loss = step(i)
    e1.record()
    torch.cuda.synchronize(dev)
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item()) / a.steps
    if eng is not None:
        eng.close()
    if hasattr(ddp, 'close'):
        ddp.close()
    return {'impl': kind, 'width': a.width, 'n_gpus': world, 'per_gpu_batch': a.batch, 'ms_per_step': ms,
            'images_per_s': a.batch * world / ms * 1e3, 'final_loss': float(loss.item()),
            'params': sum(p.numel() for p in model.parameters())}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--width', type=int, default=1)
    ap.add_argument('--batch', type=int, default=512)
    ap.add_argument('--steps', type=int, default=50)
    ap.add_argument('--warmup', type=int, default=10)
    ap.add_argument('--impl', default='both')
    ap.add_argument('--out', default='')
    a = ap.parse_args()
    rank, local, world = int(os.environ.get('RANK', 0)), int(os.environ.get('LOCAL_RANK', 0)), int(os.environ.get('WORLD_SIZE', 1))
    dev = torch.device('cuda', local)
    torch.cuda.set_device(dev)
    os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
    os.environ.setdefault('MASTER_PORT', '29577')
    dist.init_process_group('nccl', init_method='env://', world_size=world, rank=rank, device_id=dev)
    rows = []
    for kind in (('engine', 'ours', 'stock') if a.impl == 'both' else tuple(a.impl.split(','))):
        r = run(kind, a, rank, local, world, dev)
        rows.append(r)
        if rank == 0:
            print(json.dumps(r), flush=True)
    if rank == 0 and a.out:
        with open(a.out, 'w') as f:
            json.dump(rows, f, indent=1)
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
