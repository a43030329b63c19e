#!/usr/bin/env python
"""Phase timing of the fused step with the in-kernel LL all-reduce (torchrun, PDM_NET_TRACE=1)."""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault('PDM_NET_TRACE', '1')
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    rank, local, world = int(os.environ['RANK']), int(os.environ['LOCAL_RANK']), int(os.environ['WORLD_SIZE'])
    dev = torch.device('cuda', local)
    torch.cuda.set_device(dev)
    dist.init_process_group('nccl', init_method='env://', device_id=dev)
    import pytorch_distributed_mnist_b200 as pdm
    from pytorch_distributed_mnist_b200.ops import _lib
    model = pdm.DistributedDataParallel(pdm.Net().to(dev), device_ids=[local])
    opt = pdm.FusedAdam(model.parameters(), lr=1e-3)
    from pytorch_distributed_mnist_b200.ops.engine import NetEngine
    eng = NetEngine(opt.flat, opt, comm=model.comm, max_batch=512)
    B, nb, steps = 512, 64, 400
    x = torch.randint(0, 256, (nb * B, 784), dtype=torch.uint8, device=dev)
    y = torch.randint(0, 10, (nb * B,), dtype=torch.uint8, device=dev)
    eng.train_device(x, y, B, 50)
    dist.barrier()
    ms = eng.train_device(x, y, B, steps)
    lib = _lib.lib()
    lib.pdm_net_trace_read.argtypes = [ctypes.c_void_p]
    buf = np.zeros((148, 8), dtype=np.int64)
    _lib.check(lib.pdm_net_trace_read(buf.ctypes.data), 'trace')
    g = lib.pdm_net_default_ctas(B)
    t = buf[:g, :5].astype(np.float64)
    names = ['prologue', 'phase 1', 'grid barrier', 'phase 2 (+LL exchange)']
    out = ['rank {} world {}: {:.2f} us/step'.format(rank, world, ms * 1e3 / steps)]
    for i, n in enumerate(names):
        d = (t[:, i + 1] - t[:, i]) / 1.965 / 1e3
        out.append('  {:24s} mean {:6.2f} min {:6.2f} max {:6.2f} us'.format(n, d.mean(), d.min(), d.max()))
    tot = (t[:, 4] - t[:, 0]) / 1.965 / 1e3
    out.append('  {:24s} mean {:6.2f} min {:6.2f} max {:6.2f} us'.format('CTA lifetime', tot.mean(), tot.min(), tot.max()))
    for r in range(world):
        dist.barrier()
        if r == rank and r < 2:
            print('\n'.join(out), flush=True)
    eng.close()
    model.close()
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
