"""``Trainer`` -- same constructor and return values as the reference's
(ref:68-116): ``train()`` runs one epoch and returns ``(Average, Accuracy)``;
``evaluate()`` does the same over the test loader under ``no_grad``.

Two execution paths behind that surface:

* **fused** (CUDA + ``Net`` + our optimizer + our loader): the whole epoch is
  handed to the native engine -- C++ loader threads, one packed H2D copy, one
  fused kernel and one 8-byte readback per step, nothing synchronous
  (``ops.engine``).  The per-step {loss, correct} pairs come back after the
  last step and are folded exactly the way the reference folds
  ``loss.item()`` / ``Accuracy.update`` (ref:94-95), so the printed numbers
  mean the same thing.
* **generic** (CPU/gloo, the CNN, user models): the reference's loop shape
  (ref:83-95) with device-side metric accumulation instead of two
  ``.item()`` syncs per step; on CUDA the loss is the fused cross-entropy op.
"""
from __future__ import annotations

import numpy as np
import torch

from .data.loader import MNISTDataLoader
from .metrics import Accuracy, Average, DeviceMetrics
from .models.net import Net
from .ops import _lib
from .ops.loss import cross_entropy
from .optim import _FusedBase
from .utils.timing import nvtx_range


def _unwrap(model):
    return model.module if hasattr(model, 'module') and isinstance(model.module, torch.nn.Module) else model


class Trainer(object):

    def __init__(self, model, optimizer, train_loader, test_loader, device, fused='auto', ctas=0,
                 loader_workers=None, data_placement='staged'):
        self.model = model
        self.optimizer = optimizer
        self._train_loader = train_loader
        self.test_loader = test_loader
        self.device = torch.device(device)
        self.ctas = ctas
        if data_placement not in ('auto', 'staged', 'device', 'zerocopy'):
            raise ValueError('data_placement must be auto, staged, device or zerocopy')
        if data_placement == 'auto':
            # With the persistent tensor-core kernel (ops.net_tc) the in-kernel gather straight from
            # page-locked host memory is the fastest end-to-end path: producer warps stream the next
            # step's rows over PCIe while the current step computes, and the host does nothing per step.
            # Without it: the staged path needs host cores AND host memory bandwidth for its gather
            # threads (measured per step: 14.5 us on 1 GPU, 19 us on 2, 39 us on 4, 71 us on 8 GPUs of a
            # 32-core box); the zero-copy path needs neither (19 us on 1 GPU, 32 us on 8).
            import os
            world = getattr(model, 'world', 1)
            tc = False
            if self.device.type == 'cuda' and torch.cuda.is_available():
                try:
                    from .ops import net_tc
                    tc = net_tc.supported(self.device)
                except Exception:  # noqa: BLE001 - no native library: the fused path is off anyway
                    tc = False
            data_placement = 'zerocopy' if tc else (
                'staged' if (os.cpu_count() or 1) // max(1, world) >= 12 else 'zerocopy')
        self.data_placement = data_placement     # fused path only; see NetEngine.place_dataset
        self.engine = None
        self.cnn_engine = None         # native CNN step executor (ops.cnn_engine)
        self.last_train_ms = None      # device time of the last fused epoch
        self._metrics = None
        want = fused in ('auto', True, 'on')
        if want and self._fused_eligible():
            from .ops.engine import NetEngine
            workers = loader_workers if loader_workers is not None else max(1, getattr(train_loader, 'num_workers', 1))
            max_batch = max(train_loader.batch_size, test_loader.batch_size if test_loader is not None else 1)
            comm = getattr(model, 'comm', None) if getattr(model, 'uses_peer', False) else None
            self.engine = NetEngine(self.optimizer.flat, self.optimizer, comm=comm, device=self.device,
                                    max_batch=max_batch, slots=8, workers=workers)
        elif want and self._cnn_eligible():
            from .ops.cnn_engine import CnnEngine
            ddp = self.model if hasattr(self.model, 'buckets') else None
            max_batch = max(train_loader.batch_size, test_loader.batch_size if test_loader is not None else 1)
            self.cnn_engine = CnnEngine(_unwrap(self.model), self.optimizer.flat, self.optimizer, ddp=ddp,
                                        max_batch=max_batch)
        elif fused in (True, 'on'):
            raise RuntimeError('fused path requested but not available: needs a CUDA device, the native library, '
                               'the reference Net, a Fused optimizer and MNISTDataLoader inputs')
        self._prepare_train_loader()

    # The training loader is a property: binding one does the per-dataset set-up of the fused path
    # (hand the arrays to the engine, page-lock / copy them for the in-kernel gather) right away, so
    # that ``train()`` itself contains only per-epoch work -- like the reference, whose ``train()``
    # has no hidden set-up cost either (ref:77-97).
    @property
    def train_loader(self):
        return self._train_loader

    @train_loader.setter
    def train_loader(self, loader):
        self._train_loader = loader
        self._prepare_train_loader()

    def _prepare_train_loader(self):
        if self.engine is not None and isinstance(self._train_loader, MNISTDataLoader):
            self.engine.prepare(self._train_loader.images, self._train_loader.labels, self.data_placement)

    def _fused_eligible(self):
        if self.device.type != 'cuda' or not torch.cuda.is_available():
            return False
        if not isinstance(_unwrap(self.model), Net) or not isinstance(self.optimizer, _FusedBase):
            return False
        if not isinstance(self.train_loader, MNISTDataLoader):
            return False
        if self.test_loader is not None and not isinstance(self.test_loader, MNISTDataLoader):
            return False
        if self.optimizer.flat.numel != _lib.NET_FLAT or not self.optimizer.flat.params.is_cuda:
            return False
        world = getattr(self.model, 'world', 1)
        if world > 1 and not getattr(self.model, 'uses_peer', False):
            return False       # stock-collective fallback: gradients go through torch.distributed
        _lib.lib()             # on a GPU box a missing native library is an error, not a fallback
        return True

    def _cnn_eligible(self):
        from .models.cnn import CNN
        from .ops import cnn_ops
        if self.device.type != 'cuda' or not torch.cuda.is_available():
            return False
        cnn = _unwrap(self.model)
        if not isinstance(cnn, CNN) or cnn.width > 8 or not isinstance(self.optimizer, _FusedBase):
            return False
        if not self.optimizer.flat.params.is_cuda or len(self.optimizer.flat.param_list) != 8:
            return False
        world = getattr(self.model, 'world', 1)
        if world > 1 and not getattr(self.model, 'uses_peer', False):
            return False
        from .ops import gemm
        if gemm.get_precision() != 'tf32':
            return False           # the native step executor issues single-pass tf32 GEMMs; 3xTF32 goes through autograd
        return cnn_ops.available()

    def _batches(self, loader):
        """(data, target) on the device; raw uint8 when the loader can provide it (4x less PCIe
        traffic, normalisation happens in conv1)."""
        if self.cnn_engine is not None and isinstance(loader, MNISTDataLoader):
            for img, lab in loader.raw_batches():
                yield img.to(self.device, non_blocking=True), lab.to(self.device, non_blocking=True)
        else:
            for data, target in loader:
                yield data.to(self.device, non_blocking=True), target.to(self.device, non_blocking=True)

    # ------------------------------------------------------------------ train
    def train(self):
        with nvtx_range('pdm.train_epoch'):
            return self._train()

    def _train(self):
        self.model.train()
        train_loss = Average()
        train_acc = Accuracy()
        if self.engine is not None:
            loader = self.train_loader
            self.engine.prepare(loader.images, loader.labels, self.data_placement)    # no-op unless the arrays changed
            idx = loader.epoch_indices()
            steps, ms = self.engine.train(idx, loader.batch_size, ctas=self.ctas)
            self.last_train_ms = ms
            # fold the per-step {mean loss, correct} pairs the way the reference folds loss.item() and
            # Accuracy.update (ref:94-95): sample-weighted, in float64, one vector operation
            n, B = idx.numel(), loader.batch_size
            sizes = np.full(steps.shape[0], B, dtype=np.float64)
            if steps.shape[0]:
                sizes[-1] = n - (steps.shape[0] - 1) * B
            train_loss.update_sum(float(np.dot(steps[:, 0].astype(np.float64), sizes)), n)
            train_acc.update_counts(int(np.rint(steps[:, 1]).sum()), n)
            return train_loss, train_acc

        if self.cnn_engine is not None:
            self.cnn_engine.read_metrics(reset=True)
            for data, target in self._batches(self.train_loader):
                self.cnn_engine.train_step(data, target)
            loss_sum, correct, samples = self.cnn_engine.read_metrics(reset=True)
            train_loss.update_sum(loss_sum, samples)
            train_acc.update_counts(correct, samples)
            return train_loss, train_acc

        dm = self._device_metrics()
        for data, target in self.train_loader:
            data = data.to(self.device, non_blocking=True)
            target = target.to(self.device, non_blocking=True)
            output = self.model(data)
            self.optimizer.zero_grad()
            loss = cross_entropy(output, target, dm.buf)
            loss.backward()
            self.optimizer.step()
        return dm.drain(train_loss, train_acc)

    # --------------------------------------------------------------- evaluate
    @torch.no_grad()
    def evaluate(self):
        with nvtx_range('pdm.evaluate'):
            return self._evaluate()

    def _evaluate(self):
        self.model.eval()
        test_loss = Average()
        test_acc = Accuracy()
        if self.engine is not None:
            loader = self.test_loader
            self.engine.set_dataset(loader.images, loader.labels)     # evaluation is sequential: staged path
            self.engine.metrics(reset=True)
            self.engine.evaluate(loader.epoch_indices(), loader.batch_size)
            loss_sum, correct, samples = self.engine.metrics(reset=True)
            test_loss.update_sum(loss_sum, samples)
            test_acc.update_counts(correct, samples)
            return test_loss, test_acc

        if self.cnn_engine is not None:
            self.cnn_engine.read_metrics(reset=True)
            for data, target in self._batches(self.test_loader):
                self.cnn_engine.eval_step(data, target)
            loss_sum, correct, samples = self.cnn_engine.read_metrics(reset=True)
            test_loss.update_sum(loss_sum, samples)
            test_acc.update_counts(correct, samples)
            return test_loss, test_acc

        dm = self._device_metrics()
        for data, target in self.test_loader:
            data = data.to(self.device, non_blocking=True)
            target = target.to(self.device, non_blocking=True)
            output = self.model(data)
            cross_entropy(output, target, dm.buf)
        return dm.drain(test_loss, test_acc)

    def _device_metrics(self):
        if self._metrics is None:
            self._metrics = DeviceMetrics(self.device)
        self._metrics.zero_()
        return self._metrics

    def close(self):
        if self.engine is not None:
            self.engine.close()
            self.engine = None
        if self.cnn_engine is not None:
            self.cnn_engine.close()
            self.cnn_engine = None
