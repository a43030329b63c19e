#!/usr/bin/env python
"""Gradient all-reduce latency / bus-bandwidth sweep, 4 KB - 64 MB, ours vs NCCL (BASELINE.json
config 5).  Launch with torchrun on N GPUs:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29544 benchmarks/allreduce_sweep.py [--iters 200] [--out gpurun_out/allreduce_N.json]

Per size, three device-timed numbers (CUDA events on the launching stream, max over ranks):
  nccl        dist.all_reduce(SUM) on a flat fp32 tensor            (what DDP's reducer calls)
  peer        our plain all-reduce kernel (auto: one-shot < 256 KB <= two-shot)
  peer+adam   our fused all-reduce + 1/W + Adam kernel              (what our reducer launches)
  nccl+adam   dist.all_reduce + div + torch.optim.Adam(foreach) step (the stock equivalent)
  nvls(+adam) with --nvls: the experimental in-switch (multimem) kernel of parallel/nvls.py; checked
              against the peer kernel's result first, skipped with a note if the device has no multicast
Bus bandwidth = 2 (W-1)/W * bytes / time, the usual all-reduce convention.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def timed(fn, iters, warmup, device):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize(device)
    dist.barrier()
    torch.cuda.synchronize(device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize(device)
    t = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()) * 1e3          # microseconds


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--iters', type=int, default=200)
    ap.add_argument('--warmup', type=int, default=20)
    ap.add_argument('--out', default='')
    ap.add_argument('--nvls', action='store_true', help='also time the in-switch (NVLS / multimem) kernel')
    ap.add_argument('--sizes', default='', help='comma-separated byte sizes (default: the 4 KB - 64 MB sweep)')
    a = ap.parse_args()
    rank, local, world = int(os.environ['RANK']), int(os.environ['LOCAL_RANK']), int(os.environ['WORLD_SIZE'])
    dev = torch.device('cuda', local)
    torch.cuda.set_device(dev)
    dist.init_process_group('nccl', init_method='env://', device_id=dev)
    from pytorch_distributed_mnist_b200.ops import _lib
    from pytorch_distributed_mnist_b200.parallel.peer import PeerComm
    sizes = [4 << 10, 16 << 10, 31400 // 4 * 4 + 8, 64 << 10, 256 << 10, 1 << 20, 4 << 20, 16 << 20, 64 << 20]
    if a.sizes:
        sizes = [int(x) for x in a.sizes.split(',') if x]
    nmax = max(sizes) // 4
    comm = PeerComm(rank, world, dev, data_bytes=nmax * 4 * 5 + 8192)
    g, out, p, m, v = (comm.data_tensor(nmax) for _ in range(5))
    g.normal_()
    nv = None
    if a.nvls:
        from pytorch_distributed_mnist_b200.parallel.nvls import NvlsRegion
        if NvlsRegion.supported(local):
            nv = NvlsRegion(rank, world, dev, data_bytes=nmax * 4 * 5 + 8192)
            ng, nout, np_, nm, nvv = (nv.data_tensor(nmax) for _ in range(5))
            ng.copy_(g)
            # numerics first: in-switch sum vs the rank-ordered P2P sum (fp32, order may differ)
            k = min(1 << 20, nmax)
            comm.allreduce(g[:k], out[:k], average=True)
            nv.allreduce(ng[:k], nout[:k], average=True)
            torch.cuda.synchronize(dev)
            err = (out[:k] - nout[:k]).abs().max().item()
            if rank == 0:
                print('nvls vs peer all-reduce: max abs diff {:.3e}'.format(err), flush=True)
            assert err < 1e-5, 'NVLS all-reduce disagrees with the P2P kernel'
        elif rank == 0:
            print('nvls: device reports no multicast support, skipped', flush=True)
    rows = []
    for nbytes in sizes:
        n = nbytes // 4 // 4 * 4
        iters = a.iters if nbytes <= (4 << 20) else max(20, a.iters // 10)
        flat = torch.randn(n, device=dev)
        rp = torch.randn(n, device=dev, requires_grad=True)
        rp.grad = torch.zeros_like(rp)
        ropt = torch.optim.Adam([rp], lr=1e-3)
        h = _lib.Hyper(opt=_lib.PDM_OPT_ADAM, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, step=1)

        def nccl():
            dist.all_reduce(flat)

        def nccl_adam():
            dist.all_reduce(rp.grad)
            rp.grad.div_(world)
            ropt.step()

        def peer():
            comm.allreduce(g[:n], out[:n], average=True)

        def peer_adam():
            comm.allreduce_optim(g[:n], p[:n], m[:n], v[:n], h)

        def nvls_plain():
            nv.allreduce(ng[:n], nout[:n], average=True)

        def nvls_adam():
            nv.allreduce_optim(ng[:n], np_[:n], nm[:n], nvv[:n], h)

        row = {'bytes': n * 4, 'world': world}
        kinds = [('nccl', nccl), ('peer', peer), ('nccl_adam', nccl_adam), ('peer_adam', peer_adam)]
        if nv is not None:
            kinds += [('nvls', nvls_plain), ('nvls_adam', nvls_adam)]
        for name, fn in kinds:
            us = timed(fn, iters, a.warmup, dev)
            row[name + '_us'] = round(us, 2)
            row[name + '_busbw_gbs'] = round(2 * (world - 1) / world * n * 4 / (us * 1e-6) / 1e9, 2)
        comm.check_status()
        if nv is not None:
            nv.check_status()
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
    if rank == 0 and a.out:
        with open(a.out, 'w') as f:
            json.dump(rows, f, indent=1)
    if nv is not None:
        nv.close()
    comm.close()
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
