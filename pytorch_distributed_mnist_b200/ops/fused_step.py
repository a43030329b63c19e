"""Single-call Python wrappers of the fused ``Net`` kernels (``csrc/net_step.cu``).

``NetStepper`` owns the scratch / barrier words / metric accumulator one device
needs and exposes ``train_step`` / ``eval_step`` on caller-provided device
tensors.  The trainer's hot loop uses the native engine instead (no Python per
step); this class is the building block for tests, notebooks and custom loops.

Parity: ``train_step`` = ref:84-95 for one batch (SURVEY K1-K13 in one launch), ``eval_step`` =
ref:106-114 (forward, loss and top-1 only), ``forward`` = ref:126.  The hyper-parameter structs
mirror ``torch.optim.Adam`` (ref:191) and the SGD-momentum variant BASELINE.json asks for.
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib


class NetStepper(object):

    def __init__(self, device, comm=None):
        self.device = torch.device(device)
        self._lib = _lib.lib()
        self.scratch = torch.empty(_lib.NET_MAX_CTAS * _lib.NET_SCRATCH_STRIDE, dtype=torch.float32, device=self.device)
        self.sync = torch.zeros(64, dtype=torch.int32, device=self.device)
        self.metrics = torch.zeros(4, dtype=torch.float64, device=self.device)
        self.step_out = torch.zeros(2, dtype=torch.float32, device=self.device)
        self.comm = comm

    def train_step(self, x, y, params, m, v, hyper, grad_out=None, ctas=0, step_out=None, scratch=None,
                   inputs_stable=False):
        """One fused step.  x [B,784] uint8|float32, y [B] uint8|int64|int32, params/m/v flat
        PDM_NET_FLAT float32 (m, v may be None when hyper.opt allows).  Returns the
        2-float device tensor {mean loss, correct}.

        ``inputs_stable=True`` promises that no kernel queued right before this call wrote ``x`` / ``y``
        (they were copied in, or are a resident dataset): the kernel then starts loading them before it
        waits for its predecessor on the stream.  The default is the safe order."""
        x = x.contiguous()
        y = y.contiguous()
        so = step_out if step_out is not None else self.step_out
        a = _lib.NetStepArgs()
        a.x, a.x_dtype = x.data_ptr(), _lib.dtype_code(x)
        a.y, a.y_dtype = y.data_ptr(), _lib.dtype_code(y)
        a.batch, a.ctas = x.size(0), int(ctas)
        a.params = params.data_ptr()
        a.m = m.data_ptr() if m is not None else None
        a.v = v.data_ptr() if v is not None else None
        a.grad_out = grad_out.data_ptr() if grad_out is not None else None
        a.scratch = (scratch if scratch is not None else self.scratch).data_ptr()
        a.sync = self.sync.data_ptr()
        a.metrics = self.metrics.data_ptr()
        a.step_out = so.data_ptr()
        a.hyper = hyper
        a.inputs_stable = 1 if inputs_stable else 0
        if self.comm is not None and self.comm.world > 1:
            a.peers = self.comm.peers_struct(advance=True)
        _lib.check(self._lib.pdm_net_train_step(ctypes.byref(a), _lib.stream_ptr()), 'pdm_net_train_step')
        return so

    def eval_step(self, x, y, params, logits_out=None):
        x = x.contiguous()
        y = y.contiguous()
        _lib.check(self._lib.pdm_net_eval_step(
            x.data_ptr(), _lib.dtype_code(x), y.data_ptr(), _lib.dtype_code(y), x.size(0), params.data_ptr(),
            logits_out.data_ptr() if logits_out is not None else None, self.metrics.data_ptr(),
            _lib.stream_ptr()), 'pdm_net_eval_step')

    def read_metrics(self, reset=True):
        loss_sum, correct, samples, _ = self.metrics.tolist()
        if reset:
            self.metrics.zero_()
        return loss_sum, int(round(correct)), int(round(samples))


def adam_hyper(lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, step=1):
    return _lib.Hyper(opt=_lib.PDM_OPT_ADAM, lr=lr, beta1=betas[0], beta2=betas[1], eps=eps,
                      weight_decay=weight_decay, step=step)


def sgd_hyper(lr=1e-3, momentum=0.0, weight_decay=0.0, step=1):
    return _lib.Hyper(opt=_lib.PDM_OPT_SGD, lr=lr, beta1=momentum, beta2=0.0, eps=0.0,
                      weight_decay=weight_decay, step=step)


def no_opt_hyper():
    return _lib.Hyper(opt=_lib.PDM_OPT_NONE, lr=0.0, beta1=0.0, beta2=0.0, eps=0.0, weight_decay=0.0, step=1)
