#!/usr/bin/env python
"""Where a step of the persistent tensor-core kernel spends its time: per-CTA clock64 stamps
(PDM_TC_TRACE=1) averaged over steps 8..63 of a run.  Usage: python benchmarks/tc_trace.py [batch]"""
import os
import sys

os.environ.setdefault('PDM_TC_TRACE', '1')
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

CHAIN = [8, 9, 10, 1, 2, 3, 4, 11, 12, 7, 21, 27, 22]
LABELS = {8: 'M wop', 9: 'M xfull', 10: 'M fwd', 1: 'E fwd done', 2: 'E part sent', 3: 'E p1 done', 4: 'E smax+bc',
          11: 'M dl+p2', 12: 'M wgrad', 7: 'E wg done', 21: 'E db sync', 27: 'E exch', 22: 'E optim'}


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    import pytorch_distributed_mnist_b200 as pdm
    from pytorch_distributed_mnist_b200.ops import _lib
    from pytorch_distributed_mnist_b200.ops.engine import NetEngine
    dev = torch.device('cuda', 0)
    net = pdm.Net().to(dev)
    opt = pdm.FusedAdam(net.parameters(), lr=1e-3)
    eng = NetEngine(opt.flat, opt, max_batch=B)
    nb = max(2, (256 << 20) // (B * 785))
    x = torch.randint(0, 256, (nb * B, 784), dtype=torch.uint8, device=dev)
    y = torch.randint(0, 10, (nb * B,), dtype=torch.uint8, device=dev)
    eng.train_device(x, y, B, 100)
    ms = eng.train_device(x, y, B, 200)
    buf = np.zeros((13, 64, 32), dtype=np.int64)
    _lib.check(_lib.lib().pdm_net_tc_trace_read(buf.ctypes.data), 'pdm_net_tc_trace_read')
    print('batch {}: {:.2f} us/step by events (traced run)'.format(B, ms * 1e3 / 200))
    print('cycles since the CTA passed bar_wop (mean over steps 8..63)')
    print('cta ' + ' '.join('{:>11s}'.format(LABELS[k]) for k in CHAIN) + '   period  P stored   P free  publish')
    for c in range(13):
        steps = buf[c, 8:64]
        base = steps[:, 8]
        row = ' '.join('{:11.0f}'.format((steps[:, k] - base).mean()) for k in CHAIN)
        print('{:3d} {} {:8.0f} {:9.0f} {:8.0f}'.format(c, row, np.diff(steps[:, 13]).mean(), (steps[:, 18] - base).mean(),
                                                     (steps[:, 17] - base).mean()) +
              ' {:8.0f}'.format((steps[1:, 13] - steps[:-1, 22]).mean()))
    eng.close()


if __name__ == '__main__':
    main()
