"""Fused flat optimizers: ``FusedAdam`` (the reference's live optimizer,
ref:191) and ``FusedSGD`` (the commented-out one, ref:192-194, which gives the
otherwise dead ``--momentum`` / ``--wd`` flags meaning).

Both are ``torch.optim.Optimizer`` subclasses, so ``param_groups`` (which
``adjust_learning_rate`` rewrites every epoch, ref:257-261), ``state_dict()``
and ``load_state_dict()`` keep their stock layout: a checkpoint written by the
reference's ``torch.optim.Adam`` loads here and vice versa (SURVEY 5.4,
including old checkpoints whose ``step`` is a Python int).

What differs is where the state lives and who applies the update:

* ``exp_avg`` / ``exp_avg_sq`` (or ``momentum_buffer``) are views into the
  model's ``FlatState``; the whole model is updated by ONE kernel
  (``pdm_optim_step``) instead of eight ``_foreach`` launches;
* when the model is wrapped in our ``DistributedDataParallel`` with the peer
  communicator, the update is applied *inside* the all-reduce kernel of each
  gradient bucket during backward, and ``step()`` only joins the comm stream;
* for ``Net`` the trainer bypasses ``step()`` altogether: the fused step
  kernel applies the update (``ops.engine``) and advances ``step_count``.

On CPU tensors the same math runs as a handful of torch ops on the flat buffer
(the CPU/gloo configuration of BASELINE.json).
"""
from __future__ import annotations

import ctypes
import math
import warnings

import torch

from .ops import _lib
from .parallel.flat import FlatState, flat_of


def _warn_if_stale(flat):
    if getattr(flat, 'state_stale', False):
        warnings.warn('optimizer state of buckets updated by the two-shot fused all-reduce is only current on the rank '
                      'that owns each slice; call DistributedDataParallel.consolidate_optimizer_state() on every rank '
                      'before state_dict() (runner.run does this before each checkpoint)', RuntimeWarning, stacklevel=3)


class _FusedBase(torch.optim.Optimizer):
    KIND = None

    def __init__(self, params, defaults):
        params = list(params)
        if params and isinstance(params[0], dict):
            if len(params) != 1:
                raise ValueError('fused optimizers support a single parameter group')
            group = dict(params[0])
            group['params'] = list(group['params'])
            plist = group['params']
            params = [group]
        else:
            plist = params
        super().__init__(params, defaults)
        self.flat = flat_of(plist)
        if self.flat is None:
            self.flat = FlatState(plist, device=plist[0].device)
        self.flat.optimizer = self
        self.step_count = 0
        self._make_state()

    # -- state as views into the flat buffer --------------------------------
    def _make_state(self):
        raise NotImplementedError

    def _step_tensor(self):
        return torch.tensor(float(self.step_count), dtype=torch.float32)

    def zero_grad(self, set_to_none=True):
        """Gradients are views into one flat buffer: zero it (one kernel) and keep the views."""
        self.flat.zero_grads()
        self.flat.attach_grads()

    # -- hyper-parameters for the native kernels -----------------------------
    def hyper(self, step=None):
        raise NotImplementedError

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        flat = self.flat
        if flat.fused_applied:
            # a reducer kernel already applied this update during backward
            flat.fused_applied = False
            self.step_count += 1
            return loss
        self.step_count += 1
        n = flat.numel
        if flat.params.is_cuda:
            h = self.hyper(self.step_count)
            lib = _lib.lib()
            _lib.check(lib.pdm_optim_step(flat.params.data_ptr(), flat.grads.data_ptr(),
                                          flat.exp_avg.data_ptr(), flat.exp_avg_sq.data_ptr(), n,
                                          ctypes.byref(h), 1.0, _lib.stream_ptr()), 'pdm_optim_step')
        else:
            self._cpu_step(flat)
        return loss

    def _cpu_step(self, flat):
        raise NotImplementedError


class FusedAdam(_FusedBase):
    KIND = _lib.PDM_OPT_ADAM

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0):
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, amsgrad=False,
                        maximize=False, foreach=None, capturable=False, differentiable=False,
                        fused=None, decoupled_weight_decay=False)
        super().__init__(params, defaults)

    def _make_state(self):
        for i, p in enumerate(self.flat.param_list):
            self.state[p] = {'step': self._step_tensor(),
                             'exp_avg': self.flat.view_of('exp_avg', i),
                             'exp_avg_sq': self.flat.view_of('exp_avg_sq', i)}

    def hyper(self, step=None):
        g = self.param_groups[0]
        return _lib.Hyper(opt=_lib.PDM_OPT_ADAM, lr=float(g['lr']), beta1=float(g['betas'][0]),
                          beta2=float(g['betas'][1]), eps=float(g['eps']),
                          weight_decay=float(g['weight_decay']),
                          step=int(step if step is not None else self.step_count + 1))

    def _cpu_step(self, flat):
        g = self.param_groups[0]
        b1, b2 = g['betas']
        t = self.step_count
        grad = flat.grads
        if g['weight_decay'] != 0:
            grad = grad.add(flat.params, alpha=g['weight_decay'])
        flat.exp_avg.lerp_(grad, 1 - b1)
        flat.exp_avg_sq.mul_(b2).addcmul_(grad, grad, value=1 - b2)
        bc1 = 1 - b1 ** t
        bc2_sqrt = math.sqrt(1 - b2 ** t)
        denom = (flat.exp_avg_sq.sqrt() / bc2_sqrt).add_(g['eps'])
        flat.params.addcdiv_(flat.exp_avg, denom, value=-(g['lr'] / bc1))

    # -- checkpoint compatibility -------------------------------------------
    def state_dict(self):
        _warn_if_stale(self.flat)
        for p in self.flat.param_list:
            self.state[p]['step'] = self._step_tensor()
        sd = super().state_dict()
        # detach from the flat buffer so torch.save writes 31 KB, not the whole region
        for st in sd['state'].values():
            for k, v in list(st.items()):
                if torch.is_tensor(v):
                    st[k] = v.detach().clone()
        return sd

    def load_state_dict(self, state_dict):
        state = state_dict.get('state', {})
        groups = state_dict.get('param_groups', [])
        if len(groups) != 1:
            raise ValueError('expected a single param group in the checkpoint, got {}'.format(len(groups)))
        ids = groups[0]['params']
        if len(ids) != len(self.flat.param_list):
            raise ValueError('checkpoint has {} parameters, model has {}'.format(len(ids), len(self.flat.param_list)))
        steps = []
        with torch.no_grad():
            for i, pid in enumerate(ids):
                st = state.get(pid, state.get(str(pid)))
                if st is None:      # optimizer never stepped for this parameter
                    self.flat.view_of('exp_avg', i).zero_()
                    self.flat.view_of('exp_avg_sq', i).zero_()
                    continue
                self.flat.view_of('exp_avg', i).copy_(st['exp_avg'])
                self.flat.view_of('exp_avg_sq', i).copy_(st['exp_avg_sq'])
                s = st.get('step', 0)
                steps.append(int(round(float(s.item() if torch.is_tensor(s) else s))))
        self.step_count = max(steps) if steps else 0
        g = self.param_groups[0]
        for k, v in groups[0].items():      # missing keys keep this build's defaults
            if k != 'params':
                g[k] = tuple(v) if k == 'betas' else v
        self._make_state()


class FusedSGD(_FusedBase):
    KIND = _lib.PDM_OPT_SGD

    def __init__(self, params, lr=1e-3, momentum=0.0, weight_decay=0.0):
        defaults = dict(lr=lr, momentum=momentum, dampening=0, weight_decay=weight_decay,
                        nesterov=False, maximize=False, foreach=None, differentiable=False, fused=None)
        super().__init__(params, defaults)

    def _make_state(self):
        for i, p in enumerate(self.flat.param_list):
            self.state[p] = {'momentum_buffer': self.flat.view_of('exp_avg', i)}

    def hyper(self, step=None):
        g = self.param_groups[0]
        return _lib.Hyper(opt=_lib.PDM_OPT_SGD, lr=float(g['lr']), beta1=float(g['momentum']), beta2=0.0,
                          eps=0.0, weight_decay=float(g['weight_decay']),
                          step=int(step if step is not None else self.step_count + 1))

    def _cpu_step(self, flat):
        g = self.param_groups[0]
        grad = flat.grads
        if g['weight_decay'] != 0:
            grad = grad.add(flat.params, alpha=g['weight_decay'])
        if g['momentum'] != 0:
            if self.step_count == 1:
                flat.exp_avg.copy_(grad)
            else:
                flat.exp_avg.mul_(g['momentum']).add_(grad)
            grad = flat.exp_avg
        flat.params.add_(grad, alpha=-g['lr'])

    def state_dict(self):
        _warn_if_stale(self.flat)
        sd = super().state_dict()
        for st in sd['state'].values():
            for k, v in list(st.items()):
                if torch.is_tensor(v):
                    st[k] = v.detach().clone()
        sd['pdm_step_count'] = self.step_count
        return sd

    def load_state_dict(self, state_dict):
        state = state_dict.get('state', {})
        groups = state_dict.get('param_groups', [])
        ids = groups[0]['params'] if groups else []
        with torch.no_grad():
            for i, pid in enumerate(ids):
                st = state.get(pid, state.get(str(pid)))
                buf = None if st is None else st.get('momentum_buffer')
                if buf is None:
                    self.flat.view_of('exp_avg', i).zero_()
                else:
                    self.flat.view_of('exp_avg', i).copy_(buf)
        # torch's SGD seeds the buffer on the first step; a non-empty state means "already seeded"
        self.step_count = int(state_dict.get('pdm_step_count', 1 if state else 0))
        g = self.param_groups[0]
        for k, v in (groups[0].items() if groups else []):
            if k != 'params':
                g[k] = v
        self._make_state()


def build_optimizer(name, params, args):
    """``--optimizer adam`` (default, ref:191) or ``sgd`` (ref:192-194)."""
    if name == 'adam':
        return FusedAdam(params, lr=args.lr)
    if name == 'sgd':
        return FusedSGD(params, lr=args.lr, momentum=args.momentum, weight_decay=args.weight_decay)
    raise ValueError('unknown --optimizer {!r} (choose adam or sgd)'.format(name))
