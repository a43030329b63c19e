"""Device-side timing and NVTX ranges (the reference imports ``time`` and never
uses it; here every reported number is CUDA-event time, max over ranks)."""
from __future__ import annotations

import contextlib

import torch
import torch.distributed as dist


class DeviceTimer(object):
    """CUDA-event stopwatch on the current stream: ``with DeviceTimer() as t: ...; t.ms``."""

    def __init__(self, device=None):
        self.device = device
        self.ms = None

    def __enter__(self):
        torch.cuda.synchronize(self.device)
        self._e0 = torch.cuda.Event(enable_timing=True)
        self._e1 = torch.cuda.Event(enable_timing=True)
        self._e0.record()
        return self

    def __exit__(self, *exc):
        self._e1.record()
        torch.cuda.synchronize(self.device)
        self.ms = self._e0.elapsed_time(self._e1)


def max_over_ranks(value, device):
    """A per-rank scalar reduced with MAX over the job (the slowest rank defines the step)."""
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


@contextlib.contextmanager
def nvtx_range(name):
    """NVTX range when profiling is available; free otherwise."""
    pushed = False
    try:
        if torch.cuda.is_available():
            torch.cuda.nvtx.range_push(name)
            pushed = True
    except Exception:  # noqa: BLE001
        pushed = False
    try:
        yield
    finally:
        if pushed:
            torch.cuda.nvtx.range_pop()
