"""``NetEngine`` -- Python handle on the native epoch executor (``csrc/engine.cu``)
that drives the fused ``Net`` training/evaluation step (``csrc/net_step.cu``).

This is the hot path of the reference's ``Trainer.train()`` / ``evaluate()``
(ref:77-116) collapsed to: per step one packed H2D copy, one kernel launch, one
8-byte D2H copy -- all issued from a C++ loop, no Python in the step.
"""
from __future__ import annotations

import ctypes
import os

import numpy as np
import torch

from . import _lib


class NetEngine(object):

    def __init__(self, flat, optimizer, comm=None, device=None, max_batch=512, slots=8, workers=4):
        if flat.numel != _lib.NET_FLAT:
            raise ValueError('NetEngine drives the reference Net only (flat size {}, expected {})'
                             .format(flat.numel, _lib.NET_FLAT))
        self.flat = flat
        flat.layout_frozen = True      # the native engine works on raw pointers into this layout
        self.optimizer = optimizer
        self.comm = comm
        self.device = torch.device(device if device is not None else flat.params.device)
        self._lib = _lib.lib()
        self._h = ctypes.c_void_p()
        torch.cuda.set_device(self.device)
        # loader threads share the host with every other rank of the job: leave each rank a core
        # for its launch thread (the GPU boxes this was measured on have 16 cores for 8 GPUs)
        world = getattr(comm, 'world', 1) if comm is not None else 1
        workers = max(1, min(int(workers), max(2, (os.cpu_count() or 8) // max(1, world) - 1)))
        self.workers = workers
        _lib.check(self._lib.pdm_engine_create(ctypes.byref(self._h), self.device.index or 0,
                                               int(max_batch), int(slots), int(max(1, workers))),
                   'pdm_engine_create')
        self.max_batch = int(max_batch)
        comm_h = comm.handle if comm is not None and getattr(comm, 'world', 1) > 1 else None
        _lib.check(self._lib.pdm_engine_bind(self._h, flat.params.data_ptr(), flat.exp_avg.data_ptr(),
                                             flat.exp_avg_sq.data_ptr(), comm_h), 'pdm_engine_bind')
        self._dataset = None
        self._placed = None
        self._keepalive = {}
        self.launches = 0       # kernels launched by this engine (bench bookkeeping)

    # -- dataset -------------------------------------------------------------
    def set_dataset(self, images_u8, labels_u8):
        """images [N,784] uint8, labels [N] uint8 -- CPU tensors or numpy arrays, kept alive here."""
        img = images_u8.numpy() if torch.is_tensor(images_u8) else images_u8
        lab = labels_u8.numpy() if torch.is_tensor(labels_u8) else labels_u8
        img = np.ascontiguousarray(img, dtype=np.uint8).reshape(-1, _lib.NET_IN)
        lab = np.ascontiguousarray(lab, dtype=np.uint8).reshape(-1)
        if img.shape[0] != lab.shape[0]:
            raise ValueError('images/labels length mismatch')
        key = (img.ctypes.data, lab.ctypes.data, img.shape[0])
        if self._dataset is not None and self._dataset[0] == key:
            return
        self._dataset = (key, img, lab)
        _lib.check(self._lib.pdm_engine_set_dataset(self._h, img.ctypes.data, lab.ctypes.data, img.shape[0]),
                   'pdm_engine_set_dataset')
        self._placed = None

    PLACEMENTS = {'device': 0, 'zerocopy': 1}

    def place_dataset(self, placement):
        """Where the step kernel finds the training set (``set_dataset`` first):

        ``'staged'``   (default) C++ loader threads gather each batch into pinned slots, one H2D copy per step
        ``'device'``   the whole set is copied to HBM once; the kernel gathers its rows itself
        ``'zerocopy'`` the host arrays are page-locked and mapped; the kernel gathers over PCIe

        The native side keeps the last few placements, so coming back to a set that was placed before
        (``Trainer`` switching between loaders) costs a table look-up, not another page-locking."""
        if placement == 'staged':
            self._placed = None
            return
        if self._dataset is None:
            raise RuntimeError('set_dataset() must be called before place_dataset()')
        key, img, lab = self._dataset
        if self._placed == (key, placement):
            return
        _lib.check(self._lib.pdm_engine_place_dataset(self._h, img.ctypes.data, lab.ctypes.data, img.shape[0],
                                                      self.PLACEMENTS[placement]), 'pdm_engine_place_dataset')
        self._placed = (key, placement)
        self._keepalive[key] = (img, lab)       # a placement reads these arrays for as long as it is cached

    def prepare(self, images_u8, labels_u8, placement):
        """Everything a pass over this dataset needs that does not depend on the epoch."""
        self.set_dataset(images_u8, labels_u8)
        self.place_dataset(placement)

    # -- runs ------------------------------------------------------------------
    def _indices(self, indices):
        if torch.is_tensor(indices):
            indices = indices.numpy()
        return np.ascontiguousarray(indices, dtype=np.int64)

    def train(self, indices, batch, ctas=0, want_steps=True):
        """One pass of fused training steps.  Returns (per-step [n,2] float32 array of
        {mean loss, correct} or None, device milliseconds)."""
        idx = self._indices(indices)
        nsteps = (len(idx) + batch - 1) // batch
        if nsteps == 0:
            return np.zeros((0, 2), np.float32), 0.0
        out = np.empty((nsteps, 2), dtype=np.float32) if want_steps else None
        ms = ctypes.c_float(0.0)
        h = self.optimizer.hyper(self.optimizer.step_count + 1)
        torch.cuda.current_stream(self.device).synchronize()
        if getattr(self, '_placed', None) is not None:
            _lib.check(self._lib.pdm_engine_run_gather(self._h, idx.ctypes.data, len(idx), int(batch), ctypes.byref(h),
                                                       int(ctas), out.ctypes.data if out is not None else None,
                                                       ctypes.byref(ms)), 'pdm_engine_run_gather')
        else:
            _lib.check(self._lib.pdm_engine_run(self._h, idx.ctypes.data, len(idx), int(batch), 1, ctypes.byref(h),
                                                int(ctas), out.ctypes.data if out is not None else None,
                                                ctypes.byref(ms)), 'pdm_engine_run(train)')
        self.optimizer.step_count += nsteps
        self.launches += int(self._lib.pdm_engine_last_launches(self._h))
        if self.comm is not None and hasattr(self.comm, 'check_status'):
            self.comm.check_status()
        return out, float(ms.value)

    def evaluate(self, indices, batch):
        idx = self._indices(indices)
        if len(idx) == 0:
            return 0.0
        ms = ctypes.c_float(0.0)
        torch.cuda.current_stream(self.device).synchronize()
        _lib.check(self._lib.pdm_engine_run(self._h, idx.ctypes.data, len(idx), int(batch), 0, None, 0, None,
                                            ctypes.byref(ms)), 'pdm_engine_run(eval)')
        self.launches += (len(idx) + batch - 1) // batch
        return float(ms.value)

    def train_device(self, dev_images, dev_labels, batch, nsteps, ctas=0):
        """Training steps over a device-resident pool (kernel-only timing).  ``dev_images`` is
        [N, 784] uint8 (raw bytes, normalised in-kernel) or float32 (already normalised);
        ``dev_labels`` [N] uint8 / int64 / int32."""
        ms = ctypes.c_float(0.0)
        h = self.optimizer.hyper(self.optimizer.step_count + 1)
        torch.cuda.current_stream(self.device).synchronize()
        _lib.check(self._lib.pdm_engine_run_device_typed(
            self._h, dev_images.data_ptr(), _lib.dtype_code(dev_images), dev_labels.data_ptr(),
            _lib.dtype_code(dev_labels), int(dev_labels.numel()), int(batch), int(nsteps), ctypes.byref(h), int(ctas),
            ctypes.byref(ms)), 'pdm_engine_run_device')
        self.optimizer.step_count += nsteps
        self.launches += int(self._lib.pdm_engine_last_launches(self._h))
        if self.comm is not None and hasattr(self.comm, 'check_status'):
            self.comm.check_status()
        return float(ms.value)

    def metrics(self, reset=True):
        """{sum(loss*n), correct, samples} accumulated on the device since the last reset."""
        out = (ctypes.c_double * 4)()
        _lib.check(self._lib.pdm_engine_metrics(self._h, out, 1 if reset else 0), 'pdm_engine_metrics')
        return float(out[0]), int(round(out[1])), int(round(out[2]))

    # raw pointers for single-step calls (tests, the generic trainer loop)
    @property
    def scratch_ptr(self):
        return self._lib.pdm_engine_scratch(self._h)

    @property
    def sync_ptr(self):
        return self._lib.pdm_engine_sync_words(self._h)

    @property
    def metrics_ptr(self):
        return self._lib.pdm_engine_metrics_ptr(self._h)

    def close(self):
        if self._h:
            self._lib.pdm_engine_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass
