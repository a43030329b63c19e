#!/usr/bin/env python
"""Device-timed sweep of the fused Net step through the native engine (no Python per step):
per-GPU batch x CTAs.  Usage: python benchmarks/step_sweep.py [--steps 2000]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--steps', type=int, default=2000)
    ap.add_argument('--batches', default='64,512,4096')
    ap.add_argument('--ctas', default='0,16,32,64,128')
    a = ap.parse_args()
    import pytorch_distributed_mnist_b200 as pdm
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
        for c in [int(v) for v in a.ctas.split(',')]:
            eng.train_device(x, y, B, 50, ctas=c)
            ms = eng.train_device(x, y, B, a.steps, ctas=c)
            print('batch {:5d} ctas {:3d} launch {} rt {} dep {} trigger {} poll {}: {:7.2f} us/step  {:9.2f} M img/s'.format(
                B, c, os.environ.get('PDM_NET_LAUNCH', 'pdl'), os.environ.get('PDM_NET_RT', 'auto'),
                os.environ.get('PDM_NET_DEP', 'flag'), os.environ.get('PDM_NET_TRIGGER', 'start'), os.environ.get('PDM_NET_POLL_NS', '0'), ms * 1e3 / a.steps, B * a.steps / ms / 1e3), flush=True)
        eng.close()


if __name__ == '__main__':
    main()
