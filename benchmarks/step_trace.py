#!/usr/bin/env python
"""Where the fused step spends its time: per-CTA clock64 stamps (PDM_NET_TRACE=1).
Usage: PDM_NET_TRACE=1 python benchmarks/step_trace.py [--batch 512] [--ctas 0]"""
import argparse
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault('PDM_NET_TRACE', '1')

import numpy as np  # noqa: E402
import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--batch', type=int, default=512)
    ap.add_argument('--ctas', type=int, default=0)
    ap.add_argument('--steps', type=int, default=200)
    a = ap.parse_args()
    import pytorch_distributed_mnist_b200 as pdm
    from pytorch_distributed_mnist_b200.ops import _lib
    from pytorch_distributed_mnist_b200.ops.engine import NetEngine
    dev = torch.device('cuda', 0)
    net = pdm.Net().to(dev)
    opt = pdm.FusedAdam(net.parameters(), lr=1e-3)
    eng = NetEngine(opt.flat, opt, max_batch=a.batch)
    nb = 64
    x = torch.randint(0, 256, (nb * a.batch, 784), dtype=torch.uint8, device=dev)
    y = torch.randint(0, 10, (nb * a.batch,), dtype=torch.uint8, device=dev)
    ms = eng.train_device(x, y, a.batch, a.steps, ctas=a.ctas)
    lib = _lib.lib()
    lib.pdm_net_trace_read.argtypes = [ctypes.c_void_p]
    buf = np.zeros((148, 8), dtype=np.int64)
    _lib.check(lib.pdm_net_trace_read(buf.ctypes.data), 'pdm_net_trace_read')
    g = lib.pdm_net_default_ctas(a.batch) if a.ctas == 0 else a.ctas
    t = buf[:g, :5].astype(np.float64)
    ghz = 1.965
    names = ['prologue (LUT, W, first x loads)', 'phase 1 (tiles + publish)', 'grid barrier', 'phase 2 (reduce + optimizer)']
    print('batch {} ctas {}: {:.2f} us/step over {} steps'.format(a.batch, g, ms * 1e3 / a.steps, a.steps))
    for i, n in enumerate(names):
        d = (t[:, i + 1] - t[:, i]) / ghz / 1e3
        print('  {:36s} mean {:6.2f} us   min {:6.2f}   max {:6.2f}'.format(n, d.mean(), d.min(), d.max()))
    tot = (t[:, 4] - t[:, 0]) / ghz / 1e3
    print('  {:36s} mean {:6.2f} us   min {:6.2f}   max {:6.2f}'.format('CTA lifetime', tot.mean(), tot.min(), tot.max()))
    eng.close()


if __name__ == '__main__':
    main()
