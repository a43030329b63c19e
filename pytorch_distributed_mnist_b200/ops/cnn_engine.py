"""``CnnEngine`` -- Python handle on the native CNN step executor (``csrc/cnn_step.cu``).

One C call per training step: forward, loss, backward and the reducer's per-bucket fused
all-reduce + optimizer are all issued from C++ (bucket 0 overlaps the convolution backward on a
side stream).  Gradients land directly in the model's ``FlatState``; autograd is not involved.
The autograd Functions in ``ops/cnn_ops.py`` remain the general path (custom losses, hooks,
``no_sync``); this engine is what ``Trainer`` uses for the stock training loop.

Parity: one ``train_step`` is the body of the reference's loop (ref:84-95: H2D'd batch ->
``model(data)`` -> ``cross_entropy`` -> ``zero_grad``/``backward`` (DDP reducer) ->
``optimizer.step`` -> ``loss.item()`` / ``Accuracy.update``) for the CNN extension of
BASELINE.json; ``eval_step`` is ref:106-114.  Metrics stay on the device until
``read_metrics`` (one sync per epoch instead of two per step).
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib

c_void_p = ctypes.c_void_p


class CnnParams(ctypes.Structure):
    _fields_ = [(n, c_void_p) for n in ('conv1_w', 'conv1_b', 'conv2_w', 'conv2_b', 'fc1_w', 'fc1_b', 'fc2_w', 'fc2_b')]


class CnnPlan(ctypes.Structure):
    _fields_ = [('n_buckets', ctypes.c_int32), ('max_ctas', ctypes.c_int32),
                ('lo', ctypes.c_int64 * 4), ('hi', ctypes.c_int64 * 4),
                ('comm', c_void_p),
                ('params_off', ctypes.c_size_t), ('grads_off', ctypes.c_size_t),
                ('m_off', ctypes.c_size_t), ('v_off', ctypes.c_size_t),
                ('params', c_void_p), ('grads', c_void_p), ('m', c_void_p), ('v', c_void_p),
                ('hyper', _lib.Hyper), ('grad_scale', ctypes.c_float), ('nvls', c_void_p)]


@_lib.optional_decl('pdm_cnn_create')
def _d1(fn):
    fn.argtypes = [ctypes.POINTER(c_void_p), ctypes.c_int, ctypes.c_int, ctypes.c_int]


@_lib.optional_decl('pdm_cnn_eval_step')
def _d2(fn):
    fn.argtypes = [c_void_p, c_void_p, ctypes.c_int, c_void_p, ctypes.c_int, ctypes.c_int, ctypes.POINTER(CnnParams),
                   c_void_p, c_void_p, c_void_p]


@_lib.optional_decl('pdm_cnn_train_step')
def _d3(fn):
    fn.argtypes = [c_void_p, c_void_p, ctypes.c_int, c_void_p, ctypes.c_int, ctypes.c_int, ctypes.POINTER(CnnParams),
                   ctypes.POINTER(CnnParams), ctypes.c_float, ctypes.c_float, ctypes.c_uint64, c_void_p, c_void_p,
                   ctypes.POINTER(CnnPlan), c_void_p]


@_lib.optional_decl('pdm_cnn_destroy')
def _d4(fn):
    fn.argtypes = [c_void_p]


NAMES = ('conv1.weight', 'conv1.bias', 'conv2.weight', 'conv2.bias', 'fc1.weight', 'fc1.bias', 'fc2.weight', 'fc2.bias')


class CnnEngine(object):

    def __init__(self, cnn, flat, optimizer, ddp=None, max_batch=512):
        """``cnn``: a ``models.CNN`` whose parameters are views of ``flat`` (a ``FlatState``);
        ``ddp``: our ``DistributedDataParallel`` wrapper (buckets + communicator) or None."""
        self.cnn, self.flat, self.optimizer, self.ddp = cnn, flat, optimizer, ddp
        self.device = flat.params.device
        names = [n for n, _ in cnn.named_parameters()]
        if tuple(names) != NAMES or len(flat.param_list) != 8:
            raise ValueError('CnnEngine drives models.CNN only (parameters {})'.format(names))
        if cnn.width > 8:
            raise ValueError('CnnEngine supports width <= 8')
        self._lib = _lib.lib()
        self._h = c_void_p()
        torch.cuda.set_device(self.device)
        _lib.check(self._lib.pdm_cnn_create(ctypes.byref(self._h), self.device.index or 0, cnn.width, int(max_batch)),
                   'pdm_cnn_create')
        self.max_batch = int(max_batch)
        flat.layout_frozen = True      # raw pointers into the layout are captured below
        self.params = CnnParams(*[flat.view_of('params', i).data_ptr() for i in range(8)])
        self.grads = CnnParams(*[flat.view_of('grads', i).data_ptr() for i in range(8)])
        self.metrics = torch.zeros(4, dtype=torch.float64, device=self.device)
        self.loss = torch.zeros(1, dtype=torch.float32, device=self.device)
        # reducer schedule: the wrapper's buckets (reverse parameter order), or one bucket for everything
        if ddp is not None:
            ranges = [(bk['lo'], bk['hi']) for bk in ddp.buckets]
        else:
            ranges = [(0, flat.numel)]
        # the native schedule reduces bucket 0 right after the fc1 gradients: only valid when that
        # bucket holds nothing but fc1/fc2 parameters (true for the default caps); otherwise use one bucket
        if len(ranges) >= 2 and ranges[0][0] < flat.offsets[4]:
            ranges = [(0, flat.numel)]
        if len(ranges) > 4:
            raise ValueError('CnnEngine supports at most 4 buckets (got {}); raise bucket_cap_mb'.format(len(ranges)))
        self._ranges = ranges
        self.world = getattr(ddp, 'world', 1) if ddp is not None else 1
        self.comm = ddp.comm if (ddp is not None and getattr(ddp, 'uses_peer', False)) else None
        self.nvls = getattr(ddp, 'nvls', None) if ddp is not None else None
        if self.world > 1 and self.comm is None:
            raise RuntimeError('CnnEngine needs the peer communicator for multi-GPU jobs')

    def _plan(self):
        opt, flat = self.optimizer, self.flat
        plan = CnnPlan()
        plan.n_buckets = len(self._ranges)
        plan.max_ctas = 0
        for i, (lo, hi) in enumerate(self._ranges):
            plan.lo[i], plan.hi[i] = lo, hi
        if self.nvls is not None:            # flat buffers live in the multicast region: reduce in the switch
            plan.nvls = self.nvls._h
            plan.comm = None
            plan.params_off = self.nvls.offset_of(flat.params)
            plan.grads_off = self.nvls.offset_of(flat.grads)
            plan.m_off = self.nvls.offset_of(flat.exp_avg)
            plan.v_off = self.nvls.offset_of(flat.exp_avg_sq)
        elif self.comm is not None:
            plan.comm = self.comm.handle
            plan.params_off = self.comm.offset_of(flat.params)
            plan.grads_off = self.comm.offset_of(flat.grads)
            plan.m_off = self.comm.offset_of(flat.exp_avg)
            plan.v_off = self.comm.offset_of(flat.exp_avg_sq)
        else:
            plan.comm = None
        plan.params, plan.grads = flat.params.data_ptr(), flat.grads.data_ptr()
        plan.m, plan.v = flat.exp_avg.data_ptr(), flat.exp_avg_sq.data_ptr()
        plan.hyper = opt.hyper(opt.step_count + 1)
        plan.grad_scale = 1.0 / self.world
        return plan

    @staticmethod
    def _prep(x):
        x = x.reshape(x.size(0), 784)
        return x if x.is_contiguous() else x.contiguous()

    def train_step(self, x, y, seed=None):
        """Forward + loss + backward + reduce + optimizer for one batch.  ``x``: [B,1,28,28] fp32
        or [B,784] uint8 on the device; ``y``: int64 / int32 / uint8 labels.  Metrics accumulate in
        ``self.metrics`` ({loss*B, correct, B}); the mean loss is left in ``self.loss``."""
        x, y = self._prep(x), y.contiguous()
        if seed is None:
            seed = int(torch.randint(0, 2 ** 62, (1,)).item())
        plan = self._plan()
        if self.comm is not None:
            for lo, hi in self._ranges:          # buckets >= 256 KB take the two-shot kernel: state is slice-wise
                self.flat.mark_sharded(lo, hi, self.world, force=self.nvls is not None)
        train = self.cnn.training
        _lib.check(self._lib.pdm_cnn_train_step(
            self._h, x.data_ptr(), _lib.dtype_code(x), y.data_ptr(), _lib.dtype_code(y), x.size(0),
            ctypes.byref(self.params), ctypes.byref(self.grads),
            float(self.cnn.p1 if train else 0.0), float(self.cnn.p2 if train else 0.0), int(seed),
            self.metrics.data_ptr(), self.loss.data_ptr(), ctypes.byref(plan), _lib.stream_ptr()), 'pdm_cnn_train_step')
        self.optimizer.step_count += 1
        return self.loss

    def eval_step(self, x, y=None, want_logits=False):
        x = self._prep(x)
        logits = torch.empty(x.size(0), 10, device=self.device) if want_logits else None
        _lib.check(self._lib.pdm_cnn_eval_step(
            self._h, x.data_ptr(), _lib.dtype_code(x), y.contiguous().data_ptr() if y is not None else None,
            _lib.dtype_code(y) if y is not None else 0, x.size(0), ctypes.byref(self.params),
            logits.data_ptr() if logits is not None else None, self.metrics.data_ptr(), _lib.stream_ptr()),
            'pdm_cnn_eval_step')
        return logits

    def read_metrics(self, reset=True):
        loss_sum, correct, samples, _ = self.metrics.tolist()
        if reset:
            self.metrics.zero_()
        return loss_sum, int(round(correct)), int(round(samples))

    def close(self):
        if self._h:
            self._lib.pdm_cnn_destroy(self._h)
            self._h = c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass
