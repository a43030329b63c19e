#!/usr/bin/env python
"""Soak test of the cross-GPU flag protocols under injected skew (in lieu of compute-sanitizer's
racecheck, which the pool does not offer).  Launch with torchrun on N GPUs:

    PDM_SKEW_NS=20000 python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29566 benchmarks/soak.py [--steps 20000] [--out gpurun_out/soak_N.json]

PDM_SKEW_NS makes every rank pause a pseudo-random time (different per rank, CTA, warp and step) right
before it publishes data or raises a flag.  Three protocols run `--steps` steps each, in chunks of 100:
  tc        the LL push exchange inside the persistent tensor-core kernel (csrc/net_tc.cu)
  per_step  the LL exchange of the per-step kernel, pull form + flag dependency, and push form
  two_shot  the flag-ordered two-shot fused all-reduce + Adam of the reducer (csrc/comm.cu), 1 MB bucket
After every chunk all replicas must be BITWISE equal and the status word clean."""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def replicas_equal(t, world):
    others = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(others, t.contiguous())
    return all(torch.equal(o, others[0]) for o in others)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--steps', type=int, default=20000)
    ap.add_argument('--out', default='')
    a = ap.parse_args()
    rank, local, world = int(os.environ['RANK']), int(os.environ['LOCAL_RANK']), int(os.environ['WORLD_SIZE'])
    dev = torch.device('cuda', local)
    torch.cuda.set_device(dev)
    dist.init_process_group('nccl', init_method='env://', device_id=dev)
    import pytorch_distributed_mnist_b200 as pdm
    from pytorch_distributed_mnist_b200.ops import _lib
    from pytorch_distributed_mnist_b200.ops.engine import NetEngine
    B, nb, chunk = 512, 8, 100
    g = torch.Generator(device=dev).manual_seed(10 + rank)
    x = torch.randint(0, 256, (nb * B, 784), dtype=torch.uint8, device=dev, generator=g)
    y = torch.randint(0, 10, (nb * B,), dtype=torch.uint8, device=dev, generator=g)
    report = {'world': world, 'skew_ns': int(os.environ.get('PDM_SKEW_NS', '0')), 'steps': a.steps, 'protocols': {}}
    forms = [('tc', 0, {}), ('per_step_pull', int(_lib.lib().pdm_net_default_ctas(B)), {})]
    for name, ctas, _ in forms:
        torch.manual_seed(0)
        model = pdm.DistributedDataParallel(pdm.Net().to(dev), device_ids=[local])
        opt = pdm.FusedAdam(model.parameters(), lr=1e-3)
        eng = NetEngine(opt.flat, opt, comm=model.comm, device=dev, max_batch=B, slots=4, workers=1)
        ok, t0 = True, time.time()
        for _ in range(0, a.steps, chunk):
            eng.train_device(x, y, B, chunk, ctas=ctas)
            ok = ok and replicas_equal(opt.flat.params, world) and bool(torch.isfinite(opt.flat.params).all())
            model.comm.check_status()
        report['protocols'][name] = {'bitwise_equal_every_100_steps': ok, 'seconds': round(time.time() - t0, 2),
                                     'launch_form': 'one launch per 100 steps' if ctas == 0 else 'one launch per step'}
        eng.close()
        model.close()
        assert ok, name + ': replicas diverged'
    # two-shot fused all-reduce + Adam on a 1 MB bucket
    from pytorch_distributed_mnist_b200.parallel.peer import PeerComm
    n = 1 << 18
    comm = PeerComm(rank, world, dev, data_bytes=n * 4 * 4 + 8192)
    gr, p, m, v = (comm.data_tensor(n) for _ in range(4))
    p.copy_(torch.randn(n, generator=torch.Generator().manual_seed(1)).to(dev))
    ok, t0 = True, time.time()
    steps2 = max(chunk, a.steps // 4)
    for t in range(1, steps2 + 1):
        gr.copy_(torch.randn(n, device=dev, generator=g))
        h = _lib.Hyper(opt=_lib.PDM_OPT_ADAM, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, step=t)
        comm.allreduce_optim(gr, p, m, v, h, algo=2)
        if t % chunk == 0:
            torch.cuda.synchronize()
            ok = ok and replicas_equal(p, world)
            comm.check_status()
    report['protocols']['two_shot_1MB'] = {'bitwise_equal_every_100_steps': ok, 'steps': steps2, 'seconds': round(time.time() - t0, 2)}
    comm.close()
    assert ok, 'two-shot: replicas diverged'
    if rank == 0:
        print(json.dumps(report), flush=True)
        if a.out:
            with open(a.out, 'w') as f:
                json.dump(report, f, indent=1)
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
