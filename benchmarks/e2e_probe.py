#!/usr/bin/env python
"""End-to-end epoch through the engine's host path with host-side profiling on.
Usage: PDM_ENGINE_PROFILE=1 python benchmarks/e2e_probe.py [--batch 512] [--steps 2000] [--workers 6]"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--batch', type=int, default=512)
    ap.add_argument('--steps', type=int, default=2000)
    ap.add_argument('--workers', type=int, default=6)
    ap.add_argument('--slots', type=int, default=8)
    ap.add_argument('--placement', default='staged')
    a = ap.parse_args()
    import pytorch_distributed_mnist_b200 as pdm
    from pytorch_distributed_mnist_b200.ops.engine import NetEngine
    dev = torch.device('cuda', 0)
    net = pdm.Net().to(dev)
    opt = pdm.FusedAdam(net.parameters(), lr=1e-3)
    eng = NetEngine(opt.flat, opt, max_batch=a.batch, slots=a.slots, workers=a.workers)
    n = 60000
    rng = np.random.RandomState(0)
    eng.set_dataset(rng.randint(0, 256, size=(n, 784), dtype=np.uint8), rng.randint(0, 10, size=n).astype(np.uint8))
    eng.place_dataset(a.placement)
    idx = torch.from_numpy(rng.randint(0, n, size=a.steps * a.batch).astype(np.int64))
    eng.train(idx[:50 * a.batch], a.batch)
    for label, ix in (('shuffled', idx), ('sequential', torch.arange(a.steps * a.batch) % n)):
        t0 = time.perf_counter()
        _, ms = eng.train(ix, a.batch)
        wall = (time.perf_counter() - t0) * 1e3
        print('{:10s} [{}] batch {} workers {} slots {}: device {:.2f} us/step, wall {:.2f} us/step, {:.2f} M img/s'.format(
            label, a.placement, a.batch, a.workers, a.slots, ms * 1e3 / a.steps, wall * 1e3 / a.steps, a.batch * a.steps / ms / 1e3), flush=True)
    eng.close()


if __name__ == '__main__':
    main()
