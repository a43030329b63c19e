"""Python handle on the persistent tensor-core training kernel (``csrc/net_tc.cu``).

``tc_run`` executes ``nsteps`` whole training steps of the reference ``Net`` (ref:83-95: forward,
cross-entropy, backward, gradient all-reduce, Adam/SGD, metrics) in ONE launch of a 13-CTA cluster;
the model stays on chip between steps.  The trainer reaches it through ``ops.engine.NetEngine``;
this function is the building block for tests and custom loops.
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib


def supported(device=None):
    """True when the current (or given) CUDA device can launch the cluster kernel."""
    if not torch.cuda.is_available():
        return False
    dev = torch.device(device if device is not None else torch.cuda.current_device())
    index = dev.index if dev.index is not None else torch.cuda.current_device()
    return _lib.lib().pdm_net_tc_supported(int(index)) == 1


def tc_run(x, y, params, m, v, hyper, batch, nsteps=0, row_index=None, count=None, first_batch=0,
           grad_out=None, logits_out=None, metrics=None, step_out=None, peers=None, ring=None):
    """x [N,784] uint8 and y [N] uint8 (device tensors, or pinned + mapped host memory given as
    ``(data_ptr, rows)`` tuples); params/m/v flat PDM_NET_FLAT float32.  Pool mode (``row_index`` None):
    step i reads batch ``(first_batch + i) % (N // batch)``.  Plan mode: ``row_index`` int32 device tensor,
    ``count`` samples (default: all of it), last step ragged.  ``ring``: optional uint8 device tensor of
    ``_lib.TC_RING_BYTES`` bytes that enables the feeder cluster (whole-row gather into a device ring)."""
    a = _lib.NetTcArgs()
    if isinstance(x, tuple):
        a.x, rows = x
        a.y = y[0] if isinstance(y, tuple) else y.data_ptr()
    else:
        a.x, a.y, rows = x.data_ptr(), y.data_ptr(), x.size(0)
    a.pool_rows = int(rows)
    a.first_batch = int(first_batch)
    if row_index is not None:
        a.row_index = row_index.data_ptr()
        a.count = int(count if count is not None else row_index.numel())
    a.batch, a.nsteps = int(batch), int(nsteps)
    a.params = params.data_ptr()
    a.m = m.data_ptr() if m is not None else None
    a.v = v.data_ptr() if v is not None else None
    a.grad_out = grad_out.data_ptr() if grad_out is not None else None
    a.logits_out = logits_out.data_ptr() if logits_out is not None else None
    a.metrics = metrics.data_ptr() if metrics is not None else None
    a.step_out = step_out.data_ptr() if torch.is_tensor(step_out) else step_out
    a.hyper = hyper
    if peers is not None:
        a.peers = peers
    if ring is not None:
        a.ring, a.ring_bytes = ring.data_ptr(), ring.numel()
    _lib.check(_lib.lib().pdm_net_tc_run(ctypes.byref(a), _lib.stream_ptr()), 'pdm_net_tc_run')
