#!/usr/bin/env python
"""One short run of the persistent tensor-core kernel for ncu (the kernel is ONE launch; ncu replays it ~40 times,
so keep it short).  Usage: python benchmarks/profile_tc.py [steps] [batch]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402


def main():
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 512
    import pytorch_distributed_mnist_b200 as pdm
    from pytorch_distributed_mnist_b200.ops.engine import NetEngine
    dev = torch.device('cuda', 0)
    net = pdm.Net().to(dev)
    opt = pdm.FusedAdam(net.parameters(), lr=1e-3)
    eng = NetEngine(opt.flat, opt, max_batch=B)
    nb = max(2, (256 << 20) // (B * 785))
    x = torch.randint(0, 256, (nb * B, 784), dtype=torch.uint8, device=dev)
    y = torch.randint(0, 10, (nb * B,), dtype=torch.uint8, device=dev)
    ms = eng.train_device(x, y, B, steps)
    print('{} steps of batch {}: {:.2f} us/step'.format(steps, B, ms * 1e3 / steps))
    eng.close()


if __name__ == '__main__':
    main()
