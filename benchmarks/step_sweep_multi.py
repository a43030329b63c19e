#!/usr/bin/env python
"""us/step of the fused Net step with the in-kernel LL all-reduce under the launch knobs
(PDM_NET_DEP, PDM_NET_RT, PDM_NET_TRIGGER, PDM_NET_LAUNCH); run with torch.distributed.run."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    rank, local, world = int(os.environ['RANK']), int(os.environ['LOCAL_RANK']), int(os.environ['WORLD_SIZE'])
    dev = torch.device('cuda', local)
    torch.cuda.set_device(dev)
    dist.init_process_group('nccl', init_method='env://', device_id=dev)
    import pytorch_distributed_mnist_b200 as pdm
    from pytorch_distributed_mnist_b200.ops.engine import NetEngine
    model = pdm.DistributedDataParallel(pdm.Net().to(dev), device_ids=[local])
    opt = pdm.FusedAdam(model.parameters(), lr=1e-3)
    eng = NetEngine(opt.flat, opt, comm=model.comm, max_batch=512)
    B, nb, steps = 512, 512, int(os.environ.get('STEPS', 3000))
    x = torch.randint(0, 256, (nb * B, 784), dtype=torch.uint8, device=dev)
    y = torch.randint(0, 10, (nb * B,), dtype=torch.uint8, device=dev)
    eng.train_device(x, y, B, 200)
    best = 1e9
    for _ in range(3):
        torch.cuda.synchronize()
        dist.barrier()
        ms = eng.train_device(x, y, B, steps)
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        best = min(best, float(t.item()))
    mine = opt.flat.params.clone()
    others = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(others, mine)
    same = all(torch.equal(o, mine) for o in others)
    if rank == 0:
        knobs = ' '.join('{}={}'.format(k, v) for k, v in sorted(os.environ.items()) if k.startswith('PDM_NET'))
        print('world {} [{}]: {:.2f} us/step  replicas identical: {}'.format(world, knobs or 'defaults', best * 1e3 / steps, same),
              flush=True)
    eng.close()
    model.close()
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
