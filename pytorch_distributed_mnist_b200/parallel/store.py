"""``NativeStore`` -- Python face of the native TCP rendezvous store (``csrc/store.cu``).

It offers the three calls the communicator bootstrap uses (``set`` / ``get`` / ``add``) with
the same semantics as ``torch.distributed.TCPStore``: ``get`` blocks until some
rank has published the key.  ``PeerComm`` accepts either store; ``store_for_job`` builds the
native one next to the job's ``tcp://host:port`` rendezvous (port + 1 by default, override
with ``PDM_STORE_PORT``).
"""
from __future__ import annotations

import ctypes
import os

from ..ops import _lib

c_void_p = ctypes.c_void_p


@_lib.optional_decl('pdm_store_create')
def _d1(fn):
    fn.argtypes = [ctypes.POINTER(c_void_p), ctypes.c_char_p, ctypes.c_int, ctypes.c_int, ctypes.c_double]


@_lib.optional_decl('pdm_store_set')
def _d2(fn):
    fn.argtypes = [c_void_p, ctypes.c_char_p, c_void_p, ctypes.c_size_t]


@_lib.optional_decl('pdm_store_get')
def _d3(fn):
    fn.argtypes = [c_void_p, ctypes.c_char_p, c_void_p, ctypes.c_size_t, ctypes.POINTER(ctypes.c_size_t)]


@_lib.optional_decl('pdm_store_add')
def _d4(fn):
    fn.argtypes = [c_void_p, ctypes.c_char_p, ctypes.c_int64, ctypes.POINTER(ctypes.c_int64)]


@_lib.optional_decl('pdm_store_destroy')
def _d5(fn):
    fn.argtypes = [c_void_p]


class NativeStore(object):

    def __init__(self, host, port, is_server, timeout_s=300.0):
        self._lib = _lib.lib()
        self._h = c_void_p()
        _lib.check(self._lib.pdm_store_create(ctypes.byref(self._h), host.encode(), int(port),
                                              1 if is_server else 0, float(timeout_s)), 'pdm_store_create')

    def set(self, key, value):
        if isinstance(value, str):
            value = value.encode()
        value = bytes(value)
        buf = ctypes.create_string_buffer(value, len(value)) if value else None
        _lib.check(self._lib.pdm_store_set(self._h, key.encode(), buf, len(value)), 'pdm_store_set')

    def get(self, key, max_bytes=1 << 20):
        buf = ctypes.create_string_buffer(max_bytes)
        n = ctypes.c_size_t(0)
        _lib.check(self._lib.pdm_store_get(self._h, key.encode(), buf, max_bytes, ctypes.byref(n)), 'pdm_store_get')
        if n.value > max_bytes:
            raise ValueError('value of {!r} is {} bytes, larger than max_bytes={}'.format(key, n.value, max_bytes))
        return buf.raw[:n.value]

    def add(self, key, delta):
        out = ctypes.c_int64(0)
        _lib.check(self._lib.pdm_store_add(self._h, key.encode(), int(delta), ctypes.byref(out)), 'pdm_store_add')
        return out.value

    def close(self):
        if self._h:
            self._lib.pdm_store_destroy(self._h)
            self._h = c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass


def store_for_job(init_method, rank, timeout_s=300.0):
    """Native store for a job whose rendezvous URL is ``tcp://host:port``."""
    if not init_method.startswith('tcp://'):
        raise ValueError('native store needs a tcp:// init method, got {!r}'.format(init_method))
    host, port = init_method[len('tcp://'):].rsplit(':', 1)
    port = int(os.environ.get('PDM_STORE_PORT', int(port) + 1))
    return NativeStore(host, port, is_server=(rank == 0), timeout_s=timeout_s)
