#!/usr/bin/env python
"""Device-timed comparison of the two forms of the fused Net training step at several per-GPU batch
sizes: the one-launch-per-step FMA kernel (csrc/net_step.cu) and the persistent tensor-core cluster
kernel (csrc/net_tc.cu).  Both through NetEngine.train_device (resident uint8 pool larger than L2).
Usage: python benchmarks/tc_sweep.py [--steps 2000] [--batches 64,512,4096,16384]"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--steps', type=int, default=2000)
    ap.add_argument('--batches', default='64,512,4096,16384')
    a = ap.parse_args()
    import pytorch_distributed_mnist_b200 as pdm
    from pytorch_distributed_mnist_b200.ops import _lib
    from pytorch_distributed_mnist_b200.ops.engine import NetEngine
    dev = torch.device('cuda', 0)
    for B in [int(x) for x in a.batches.split(',')]:
        torch.manual_seed(0)
        net = pdm.Net().to(dev)
        opt = pdm.FusedAdam(net.parameters(), lr=1e-3)
        eng = NetEngine(opt.flat, opt, max_batch=B)
        nb = max(2, (256 << 20) // (B * 785))
        x = torch.randint(0, 256, (nb * B, 784), dtype=torch.uint8, device=dev)
        y = torch.randint(0, 10, (nb * B,), dtype=torch.uint8, device=dev)
        steps = max(50, a.steps * 512 // max(512, B))
        row = {'batch': B, 'steps': steps}
        # an explicit CTA count selects the per-step FMA kernel, 0 the persistent tensor-core kernel
        for name, ctas in (('tc_persistent', 0), ('fma_per_step', int(_lib.lib().pdm_net_default_ctas(B)))):
            eng.train_device(x, y, B, 50, ctas=ctas)
            l0 = eng.launches
            ms = eng.train_device(x, y, B, steps, ctas=ctas)
            row[name + '_us_per_step'] = round(ms * 1e3 / steps, 3)
            row[name + '_Mimg_s'] = round(B * steps / ms / 1e3, 2)
            row[name + '_launches'] = eng.launches - l0
        print(json.dumps(row), flush=True)
        eng.close()


if __name__ == '__main__':
    main()
