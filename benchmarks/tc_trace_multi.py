#!/usr/bin/env python
"""Phase trace of the persistent tensor-core kernel with the in-kernel LL all-reduce, one rank per GPU:
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 benchmarks/tc_trace_multi.py
Prints, for ranks 0 and N-1, the cycles each CTA spends between the stamps around the exchange
(mean over steps 8..63) and the step period."""
import os
import sys

os.environ.setdefault('PDM_TC_TRACE', '1')
os.environ.setdefault('PDM_NET_TC_MAX_WORLD', '8')
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

SEG = [(13, 1, 'publ>fwd'), (1, 4, 'fwd>smax'), (4, 7, 'smax>wg'), (7, 21, 'wg>db'), (21, 24, 'pushed'), (24, 25, 'polled'),
       (25, 26, 'cta sync'), (26, 27, 'summed'), (27, 22, 'optim')]


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    rank, world = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
    local = int(os.environ.get('LOCAL_RANK', rank))
    torch.cuda.set_device(local)
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    import pytorch_distributed_mnist_b200 as pdm
    from pytorch_distributed_mnist_b200.ops import _lib
    from pytorch_distributed_mnist_b200.ops.engine import NetEngine
    dev = torch.device('cuda', local)
    model = pdm.DistributedDataParallel(pdm.Net().to(dev), device_ids=[local])
    opt = pdm.FusedAdam(model.parameters(), lr=1e-3)
    eng = NetEngine(model.flat, opt, comm=model.comm, max_batch=B)
    nb = max(2, (256 << 20) // (B * 785))
    g = torch.Generator(device='cpu').manual_seed(rank)
    x = torch.randint(0, 256, (nb * B, 784), dtype=torch.uint8, generator=g).to(dev)
    y = torch.randint(0, 10, (nb * B,), dtype=torch.uint8, generator=g).to(dev)
    eng.train_device(x, y, B, 100)
    dist.barrier()
    torch.cuda.synchronize()
    ms = eng.train_device(x, y, B, 200)
    t = torch.tensor([ms], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    buf = np.zeros((13, 64, 32), dtype=np.int64)
    _lib.check(_lib.lib().pdm_net_tc_trace_read(buf.ctypes.data), 'pdm_net_tc_trace_read')
    for r in (0, world - 1):
        dist.barrier()
        if rank != r:
            continue
        print('rank {} of {}, batch {}: {:.2f} us/step (max over ranks, traced run), launches {}'.format(
            r, world, B, t.item() * 1e3 / 200, eng.launches), flush=True)
        print('cta ' + ' '.join('{:>9s}'.format(s[2]) for s in SEG) + '    period', flush=True)
        for c in (0, 6, 12):
            st = buf[c, 8:64]
            row = ' '.join('{:9.0f}'.format((st[:, b] - st[:, a]).mean()) for a, b, _ in SEG[1:])
            first = (st[1:, 1] - st[:-1, 13]).mean()
            print('{:3d} {:9.0f} {} {:9.0f}'.format(c, first, row, np.diff(st[:, 13]).mean()), flush=True)
        pol = buf[:, 8:64, 25] - buf[:, 8:64, 24]
        print('polled: min {:.0f} median {:.0f} p90 {:.0f} max {:.0f} cycles over CTAs x steps'.format(
            pol.min(), np.median(pol), np.percentile(pol, 90), pol.max()), flush=True)
    dist.barrier()
    eng.close()
    model.close()
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
