#!/usr/bin/env python
"""How long does the per-step input copy take?  Pinned host -> device, CUDA-event timed."""
import torch

dev = torch.device('cuda', 0)
for nbytes in (8, 401920, 1609728, 16 << 20):
    h = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    d = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    for _ in range(20):
        d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 500
    e0.record()
    for _ in range(n):
        d.copy_(h, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / n
    print('H2D {:9d} B: {:8.2f} us/copy  {:7.2f} GB/s'.format(nbytes, us, nbytes / us / 1e3), flush=True)
    e0.record()
    for _ in range(n):
        h.copy_(d, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / n
    print('D2H {:9d} B: {:8.2f} us/copy  {:7.2f} GB/s'.format(nbytes, us, nbytes / us / 1e3), flush=True)
