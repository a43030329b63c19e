"""``NvlsRegion`` -- symmetric memory bound to an NVSwitch multicast object (``csrc/nvls.cu``).

EXPERIMENTAL, opt-in.  Written and compiled in round 1 but not yet run on hardware; nothing in
the training paths uses it.  The P2P two-shot kernel of ``PeerComm`` stays the large-bucket path
until this one has been measured against it (``benchmarks/allreduce_sweep.py --nvls``).

Role: NCCL reduces large messages inside the switch (NVLS) when the fabric
allows it.  With a multicast mapping, ``multimem.ld_reduce`` returns the sum of an address over all
replicas and ``multimem.st`` writes all replicas, so a fused all-reduce + optimizer moves ~1/W of
the bucket per GPU in each direction instead of (W-1)/W twice.

Bootstrap: the multicast object is created by rank 0 and shared as a POSIX file descriptor.  A
descriptor cannot travel through the TCP store, so rank 0 listens on an abstract Unix socket
(name published through the store) and hands the descriptor to each peer with ``SCM_RIGHTS``
(``socket.send_fds``); everything else (add device -> barrier -> bind -> barrier) is ordered with
store counters, exactly like ``PeerComm``'s handle exchange.
"""
from __future__ import annotations

import ctypes
import itertools
import os
import socket
import time

import torch

from ..ops import _lib
from .peer import _CudaBlock, default_store

c_void_p = ctypes.c_void_p
_counter = itertools.count()


@_lib.optional_decl('pdm_nvls_supported')
def _d0(fn):
    fn.argtypes = [ctypes.c_int]


@_lib.optional_decl('pdm_nvls_create')
def _d1(fn):
    fn.argtypes = [ctypes.POINTER(c_void_p), ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_size_t]


@_lib.optional_decl('pdm_nvls_export_fd')
def _d2(fn):
    fn.argtypes = [c_void_p, ctypes.POINTER(ctypes.c_int)]


@_lib.optional_decl('pdm_nvls_import_fd')
def _d3(fn):
    fn.argtypes = [c_void_p, ctypes.c_int]


@_lib.optional_decl('pdm_nvls_add_device')
def _d4(fn):
    fn.argtypes = [c_void_p]


@_lib.optional_decl('pdm_nvls_bind')
def _d5(fn):
    fn.argtypes = [c_void_p]


@_lib.optional_decl('pdm_nvls_data')
def _d6(fn):
    fn.argtypes = [c_void_p]
    fn.restype = c_void_p


@_lib.optional_decl('pdm_nvls_data_bytes')
def _d7(fn):
    fn.argtypes = [c_void_p]
    fn.restype = ctypes.c_size_t


@_lib.optional_decl('pdm_nvls_set_timeout')
def _d8(fn):
    fn.argtypes = [c_void_p, ctypes.c_double]
    fn.restype = None


@_lib.optional_decl('pdm_nvls_status')
def _d9(fn):
    fn.argtypes = [c_void_p, ctypes.POINTER(ctypes.c_int)]


@_lib.optional_decl('pdm_nvls_allreduce_optim')
def _d10(fn):
    fn.argtypes = [c_void_p, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_size_t,
                   ctypes.POINTER(_lib.Hyper), ctypes.c_float, c_void_p]


@_lib.optional_decl('pdm_nvls_destroy')
def _d11(fn):
    fn.argtypes = [c_void_p]


def _wait_count(store, key, target, timeout_s):
    store.add(key, 1)
    t0 = time.time()
    while int(store.add(key, 0)) < target:
        if time.time() - t0 > timeout_s:
            raise RuntimeError('NVLS bootstrap: only {} of {} ranks reached "{}"'.format(int(store.add(key, 0)), target, key))
        time.sleep(0.001)


def share_fd(store, tag, rank, world, fd=None, timeout_s=60.0):
    """Rank 0 passes ``fd`` to every other rank over an abstract Unix socket; returns the local fd."""
    if rank == 0:
        name = '\0{}-{}'.format(tag.replace('/', '-'), os.getpid())        # abstract namespace: no file to clean up
        srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        srv.bind(name)
        srv.listen(world)
        srv.settimeout(timeout_s)
        store.set('{}/sock'.format(tag), name[1:].encode())
        try:
            for _ in range(world - 1):
                conn, _addr = srv.accept()
                with conn:
                    socket.send_fds(conn, [b'fd'], [fd])
                    conn.recv(1)             # peer confirms it holds its own copy before we close ours
        finally:
            srv.close()
        return fd
    path = '\0' + bytes(store.get('{}/sock'.format(tag))).decode()
    cli = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    cli.settimeout(timeout_s)
    cli.connect(path)
    with cli:
        _msg, fds, _flags, _addr = socket.recv_fds(cli, 16, 1)
        cli.send(b'k')
    if len(fds) != 1:
        raise RuntimeError('NVLS bootstrap: expected one descriptor from rank 0, got {}'.format(len(fds)))
    return fds[0]


class NvlsRegion(object):
    """Symmetric region whose ``data_tensor`` slices can be all-reduced in the switch."""

    def __init__(self, rank, world, device, data_bytes, store=None, timeout_s=30.0, tag=None):
        self.rank, self.world = int(rank), int(world)
        self.device = torch.device(device)
        self._lib = _lib.lib()
        if not hasattr(self._lib, 'pdm_nvls_create'):
            raise RuntimeError('this build of the native library has no NVLS support')
        self._h = c_void_p()
        torch.cuda.set_device(self.device)
        store = store if store is not None else default_store()
        tag = tag if tag is not None else 'pdm/nvls{}'.format(next(_counter))
        try:
            _lib.check(self._lib.pdm_nvls_create(ctypes.byref(self._h), self.rank, self.world, self.device.index or 0,
                                                 int(data_bytes)), 'pdm_nvls_create')
            if self.rank == 0:
                fd = ctypes.c_int(-1)
                _lib.check(self._lib.pdm_nvls_export_fd(self._h, ctypes.byref(fd)), 'pdm_nvls_export_fd')
                share_fd(store, tag, 0, self.world, fd.value)
                os.close(fd.value)
            else:
                fd = share_fd(store, tag, self.rank, self.world)
                try:
                    _lib.check(self._lib.pdm_nvls_import_fd(self._h, fd), 'pdm_nvls_import_fd')
                finally:
                    os.close(fd)
            _lib.check(self._lib.pdm_nvls_add_device(self._h), 'pdm_nvls_add_device')
            _wait_count(store, '{}/added'.format(tag), self.world, 60.0)
            _lib.check(self._lib.pdm_nvls_bind(self._h), 'pdm_nvls_bind')
            _wait_count(store, '{}/bound'.format(tag), self.world, 60.0)
        except Exception:
            self.close()
            raise
        self._lib.pdm_nvls_set_timeout(self._h, float(timeout_s))
        self.data_bytes = int(self._lib.pdm_nvls_data_bytes(self._h))
        self._data_ptr = int(self._lib.pdm_nvls_data(self._h))
        self._cursor = 0
        self._block = _CudaBlock(self._data_ptr, self.data_bytes, self)
        self._bytes = torch.as_tensor(self._block, device=self.device)

    @staticmethod
    def supported(device=0):
        if not torch.cuda.is_available() or not _lib.available():
            return False
        lib = _lib.lib()
        return hasattr(lib, 'pdm_nvls_supported') and lib.pdm_nvls_supported(int(device)) == 1

    def data_tensor(self, numel_f32):
        nbytes = (int(numel_f32) * 4 + 255) // 256 * 256
        if self._cursor + nbytes > self.data_bytes:
            raise RuntimeError('NVLS region exhausted: need {} more bytes, {} left'.format(nbytes, self.data_bytes - self._cursor))
        lo = self._cursor
        self._cursor += nbytes
        return self._bytes[lo:lo + int(numel_f32) * 4].view(torch.float32)

    def offset_of(self, tensor):
        off = tensor.data_ptr() - self._data_ptr
        if off < 0 or off >= self.data_bytes:
            raise ValueError('tensor is not inside this NVLS region')
        return off

    def allreduce_optim(self, grads, out, exp_avg, exp_avg_sq, hyper, grad_scale=None, stream=None):
        scale = (1.0 / self.world) if grad_scale is None else grad_scale
        _lib.check(self._lib.pdm_nvls_allreduce_optim(
            self._h, self.offset_of(grads), self.offset_of(out),
            self.offset_of(exp_avg) if exp_avg is not None else 0,
            self.offset_of(exp_avg_sq) if exp_avg_sq is not None else 0,
            grads.numel(), ctypes.byref(hyper), scale, _lib.stream_ptr(stream)), 'pdm_nvls_allreduce_optim')

    def allreduce(self, src, dst, average=True, stream=None):
        """Plain all-reduce (mean by default) of ``src`` into ``dst`` on every rank (may alias)."""
        from ..ops.fused_step import no_opt_hyper
        self.allreduce_optim(src, dst, None, None, no_opt_hyper(), grad_scale=(1.0 / self.world) if average else 1.0,
                             stream=stream)

    @property
    def handle(self):
        return self._h

    def check_status(self):
        st = ctypes.c_int(0)
        _lib.check(self._lib.pdm_nvls_status(self._h, ctypes.byref(st)), 'pdm_nvls_status')
        if st.value:
            raise RuntimeError('NVLS collective timed out waiting for the other ranks')

    def close(self):
        if self._h:
            self._lib.pdm_nvls_destroy(self._h)
            self._h = c_void_p()
