"""Flat parameter / gradient / optimizer-state storage.

The reference keeps every tensor separate: DDP copies each gradient into a flat
bucket, all-reduces, copies back (4 launches), then Adam walks
the parameter list again (8 launches, K11).  Here a model owns ONE fp32 buffer
laid out as ``[params | grads | exp_avg | exp_avg_sq]``; every ``nn.Parameter``,
its ``.grad`` and its optimizer state are views into it.  Consequences:

* a gradient bucket is a contiguous slice -- no gather/scatter copies;
* the optimizer is one kernel over a slice, fused with the all-reduce;
* when the buffer lives in the communicator's symmetric region, peers read
  gradients and write parameters in place over NVLink (``parallel.peer``).

Layout rule: parameters in ``module.parameters()`` order, each starting on a
4-float (16-byte) boundary; for ``Net`` this gives exactly the
``PDM_NET_FLAT`` layout of ``include/pdmnist.h`` (weight at 0, bias at 7840).
"""
from __future__ import annotations

import torch


def _round_up(x, m):
    return (x + m - 1) // m * m


class FlatState(object):
    SECTIONS = ('params', 'grads', 'exp_avg', 'exp_avg_sq')

    def __init__(self, params, device=None, comm=None):
        params = [p for p in params]
        if not params:
            raise ValueError('FlatState needs at least one parameter')
        for p in params:
            if p.dtype != torch.float32:
                raise TypeError('FlatState supports fp32 parameters only (got {})'.format(p.dtype))
        self.param_list = params
        self.device = torch.device(device) if device is not None else params[0].device
        self.offsets = []
        off = 0
        for p in params:
            self.offsets.append(off)
            off = _round_up(off + p.numel(), 4)
        self.numel = off                              # floats, multiple of 4
        self.section = _round_up(self.numel, 64)      # section stride (256-byte aligned)
        self.comm = comm
        total = self.section * len(self.SECTIONS)
        if comm is not None:
            self.storage = comm.data_tensor(total)    # symmetric memory, zeroed at creation
        else:
            self.storage = torch.zeros(total, dtype=torch.float32, device=self.device)
        for i, name in enumerate(self.SECTIONS):
            setattr(self, name, self.storage[i * self.section:i * self.section + self.numel])
        self.optimizer = None          # set by FusedOptimizer
        self.fused_applied = False     # a reducer already applied this step's update
        with torch.no_grad():
            for p, o in zip(params, self.offsets):
                view = self.params[o:o + p.numel()].view(p.shape)
                view.copy_(p.data)
                p.data = view
                p._pdm_flat = self
        self.attach_grads()

    # byte offsets of the sections inside the storage (for the native comm calls)
    def byte_offset(self, name):
        return self.SECTIONS.index(name) * self.section * 4

    def view_of(self, name, index):
        p, o = self.param_list[index], self.offsets[index]
        return getattr(self, name)[o:o + p.numel()].view(p.shape)

    def attach_grads(self):
        """Point every ``p.grad`` at its slice of the flat gradient buffer."""
        for i, p in enumerate(self.param_list):
            if p.requires_grad:
                p.grad = self.view_of('grads', i)

    def zero_grads(self):
        self.grads.zero_()

    def range_of(self, first, last):
        """Float range [lo, hi) covering parameters first..last (inclusive), 4-aligned."""
        lo = self.offsets[first]
        hi = _round_up(self.offsets[last] + self.param_list[last].numel(), 4)
        return lo, hi


def flat_of(params):
    """The FlatState shared by all ``params`` (in order), or None."""
    params = list(params)
    flats = {id(getattr(p, '_pdm_flat', None)) for p in params}
    if len(flats) != 1:
        return None
    flat = getattr(params[0], '_pdm_flat', None)
    if flat is None or len(flat.param_list) != len(params):
        return None
    if any(a is not b for a, b in zip(flat.param_list, params)):
        return None
    return flat
