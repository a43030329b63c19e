#!/usr/bin/env python
"""Short single-GPU driver for ncu: launches the fused Net step kernel a few times.
Usage: python benchmarks/profile_step.py [--batch 512] [--steps 12] [--ctas 0] [--dtype u8|f32]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--batch', type=int, default=512)
    ap.add_argument('--steps', type=int, default=12)
    ap.add_argument('--ctas', type=int, default=0)
    ap.add_argument('--dtype', default='u8')
    a = ap.parse_args()
    import pytorch_distributed_mnist_b200 as pdm
    from pytorch_distributed_mnist_b200.ops.fused_step import NetStepper
    dev = torch.device('cuda', 0)
    torch.manual_seed(0)
    net = pdm.Net().to(dev)
    opt = pdm.FusedAdam(net.parameters(), lr=1e-3)
    st = NetStepper(dev)
    nb = 64
    x = torch.randint(0, 256, (nb, a.batch, 784), dtype=torch.uint8, device=dev)
    if a.dtype == 'f32':
        x = x.float().div(255).sub(0.1307).div(0.3081)
    y = torch.randint(0, 10, (nb, a.batch), dtype=torch.uint8, device=dev)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for t in range(1, 4):
        st.train_step(x[t % nb], y[t % nb], opt.flat.params, opt.flat.exp_avg, opt.flat.exp_avg_sq, opt.hyper(t), ctas=a.ctas)
    torch.cuda.synchronize()
    ev0.record()
    for t in range(4, 4 + a.steps):
        st.train_step(x[t % nb], y[t % nb], opt.flat.params, opt.flat.exp_avg, opt.flat.exp_avg_sq, opt.hyper(t), ctas=a.ctas)
    ev1.record()
    torch.cuda.synchronize()
    print('batch {} ctas {} dtype {}: {:.2f} us/step over {} python-launched steps'.format(
        a.batch, a.ctas, a.dtype, ev0.elapsed_time(ev1) * 1e3 / a.steps, a.steps))


if __name__ == '__main__':
    main()
