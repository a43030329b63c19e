"""ctypes binding of ``libpdmnist_b200.so`` (C ABI in ``include/pdmnist.h``).

The library is loaded from ``pytorch_distributed_mnist_b200/_native/`` (built
in-tree by ``ops/_build.py``).  Policy on a missing library:

* on a box with a CUDA device, ``lib()`` raises -- the GPU paths must fail
  loudly rather than fall back to eager PyTorch;
* on a CPU-only box ``available()`` is simply False and the CPU paths never
  ask for it (host-only pieces -- store, loader threads -- still load fine
  because the library links the CUDA runtime statically).
"""
from __future__ import annotations

import ctypes
import os
import threading

from . import _build

_lock = threading.Lock()
_lib = None
_load_error = None

c_void_p = ctypes.c_void_p
c_int = ctypes.c_int
c_float = ctypes.c_float
c_size_t = ctypes.c_size_t

PDM_U8, PDM_F32, PDM_I64, PDM_I32 = 0, 1, 2, 3
PDM_OPT_ADAM, PDM_OPT_SGD, PDM_OPT_NONE = 0, 1, 2
PDM_MAX_RANKS = 8
NET_IN, NET_OUT, NET_PARAMS, NET_FLAT = 784, 10, 7850, 7852
NET_SCRATCH_STRIDE = 7856
NET_MAX_CTAS = 148
NET_STATIC_MAX_CTAS = 64


class Hyper(ctypes.Structure):
    _fields_ = [('opt', ctypes.c_int32), ('lr', c_float), ('beta1', c_float), ('beta2', c_float),
                ('eps', c_float), ('weight_decay', c_float), ('step', ctypes.c_int64)]


class Peers(ctypes.Structure):
    _fields_ = [('world', ctypes.c_int32), ('rank', ctypes.c_int32), ('seq', ctypes.c_uint32),
                ('pad_', ctypes.c_uint32), ('inbox', c_void_p * PDM_MAX_RANKS),
                ('status', c_void_p), ('timeout_ns', ctypes.c_uint64)]


class NetStepArgs(ctypes.Structure):
    _fields_ = [('x', c_void_p), ('x_dtype', ctypes.c_int32),
                ('y', c_void_p), ('y_dtype', ctypes.c_int32),
                ('batch', ctypes.c_int32), ('ctas', ctypes.c_int32),
                ('params', c_void_p), ('m', c_void_p), ('v', c_void_p),
                ('grad_out', c_void_p), ('scratch', c_void_p), ('sync', c_void_p),
                ('metrics', c_void_p), ('step_out', c_void_p),
                ('hyper', Hyper), ('peers', Peers), ('row_index', c_void_p),
                ('flag_dependency', ctypes.c_int32), ('done_before', ctypes.c_uint32),
                ('inputs_stable', ctypes.c_int32)]


class NetTcArgs(ctypes.Structure):
    _fields_ = [('x', c_void_p), ('y', c_void_p), ('row_index', c_void_p),
                ('pool_rows', ctypes.c_int64), ('first_batch', ctypes.c_int64), ('count', ctypes.c_int64),
                ('batch', ctypes.c_int32), ('nsteps', ctypes.c_int64),
                ('params', c_void_p), ('m', c_void_p), ('v', c_void_p),
                ('grad_out', c_void_p), ('logits_out', c_void_p), ('metrics', c_void_p), ('step_out', c_void_p),
                ('hyper', Hyper), ('peers', Peers), ('ring', c_void_p), ('ring_bytes', c_size_t)]


TC_RING_BYTES = 4 * (512 * 784 + 512) + 256


class NativeError(RuntimeError):
    pass


def _declare(lib):
    lib.pdm_last_error.restype = ctypes.c_char_p
    lib.pdm_last_error.argtypes = []
    lib.pdm_version.restype = c_int
    lib.pdm_device_ok.argtypes = [c_int]
    lib.pdm_net_train_step.argtypes = [ctypes.POINTER(NetStepArgs), c_void_p]
    lib.pdm_net_default_ctas.argtypes = [c_int]
    if hasattr(lib, 'pdm_net_train_step_static'):
        lib.pdm_net_train_step_static.argtypes = [ctypes.POINTER(NetStepArgs), c_void_p]
    lib.pdm_net_tc_supported.argtypes = [c_int]
    lib.pdm_net_tc_run.argtypes = [ctypes.POINTER(NetTcArgs), c_void_p]
    lib.pdm_net_tc_trace_read.argtypes = [c_void_p]
    lib.pdm_net_eval_step.argtypes = [c_void_p, c_int, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]
    lib.pdm_net_forward.argtypes = [c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p]
    lib.pdm_optim_step.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_size_t,
                                   ctypes.POINTER(Hyper), c_float, c_void_p]
    lib.pdm_cross_entropy.argtypes = [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]
    # communicator
    lib.pdm_comm_create.argtypes = [ctypes.POINTER(c_void_p), c_int, c_int, c_int, c_size_t]
    lib.pdm_comm_handle.argtypes = [c_void_p, c_void_p]
    lib.pdm_comm_connect.argtypes = [c_void_p, c_void_p]
    lib.pdm_comm_data.restype = c_void_p
    lib.pdm_comm_data.argtypes = [c_void_p]
    lib.pdm_comm_data_bytes.restype = c_size_t
    lib.pdm_comm_data_bytes.argtypes = [c_void_p]
    lib.pdm_comm_set_timeout.argtypes = [c_void_p, ctypes.c_double]
    lib.pdm_comm_peers.argtypes = [c_void_p, ctypes.POINTER(Peers), c_int]
    lib.pdm_comm_status.argtypes = [c_void_p, ctypes.POINTER(c_int)]
    lib.pdm_comm_barrier.argtypes = [c_void_p, c_void_p]
    lib.pdm_comm_broadcast.argtypes = [c_void_p, c_size_t, c_size_t, c_int, c_void_p]
    lib.pdm_comm_allreduce_optim.argtypes = [c_void_p, c_size_t, c_size_t, c_size_t, c_size_t, c_size_t,
                                             ctypes.POINTER(Hyper), c_float, c_int, c_int, c_void_p]
    lib.pdm_comm_destroy.argtypes = [c_void_p]
    # engine
    lib.pdm_engine_create.argtypes = [ctypes.POINTER(c_void_p), c_int, c_int, c_int, c_int]
    lib.pdm_engine_bind.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]
    lib.pdm_engine_set_dataset.argtypes = [c_void_p, c_void_p, c_void_p, ctypes.c_int64]
    lib.pdm_engine_run.argtypes = [c_void_p, c_void_p, ctypes.c_int64, c_int, c_int, ctypes.POINTER(Hyper), c_int,
                                   c_void_p, ctypes.POINTER(c_float)]
    lib.pdm_engine_run_device.argtypes = [c_void_p, c_void_p, c_void_p, ctypes.c_int64, c_int, ctypes.c_int64,
                                          ctypes.POINTER(Hyper), c_int, ctypes.POINTER(c_float)]
    lib.pdm_engine_run_device_typed.argtypes = [c_void_p, c_void_p, c_int, c_void_p, c_int, ctypes.c_int64, c_int,
                                                ctypes.c_int64, ctypes.POINTER(Hyper), c_int, ctypes.POINTER(c_float)]
    lib.pdm_engine_metrics.argtypes = [c_void_p, ctypes.POINTER(ctypes.c_double), c_int]
    lib.pdm_engine_last_launches.argtypes = [c_void_p]
    lib.pdm_engine_last_launches.restype = ctypes.c_int64
    lib.pdm_engine_place_dataset.argtypes = [c_void_p, c_void_p, c_void_p, ctypes.c_int64, c_int]
    lib.pdm_engine_run_gather.argtypes = [c_void_p, c_void_p, ctypes.c_int64, c_int, ctypes.POINTER(Hyper), c_int,
                                          c_void_p, ctypes.POINTER(c_float)]
    for nm in ('pdm_engine_scratch', 'pdm_engine_sync_words', 'pdm_engine_metrics_ptr'):
        getattr(lib, nm).restype = c_void_p
        getattr(lib, nm).argtypes = [c_void_p]
    lib.pdm_engine_destroy.argtypes = [c_void_p]
    for name, fn in _OPTIONAL_DECLS.items():
        if hasattr(lib, name):
            fn(getattr(lib, name))


_OPTIONAL_DECLS = {}


def optional_decl(name):
    """Register argtypes for a symbol that newer builds of the library export."""
    def deco(fn):
        _OPTIONAL_DECLS[name] = fn
        if _lib is not None and hasattr(_lib, name):
            fn(getattr(_lib, name))
        return fn
    return deco


def lib_path():
    return _build.LIB_PATH


def load(build_if_missing=True):
    """Load (building first if needed and possible) and return the CDLL."""
    global _lib, _load_error
    with _lock:
        if _lib is not None:
            return _lib
        path = lib_path()
        try:
            if build_if_missing and (not os.path.isfile(path) or os.environ.get('PDM_REBUILD')):
                _build.build()
            lib = ctypes.CDLL(path, mode=ctypes.RTLD_GLOBAL)
            _declare(lib)
            _lib = lib
            return _lib
        except Exception as e:  # noqa: BLE001 - surfaced by lib()
            _load_error = e
            raise


def available():
    try:
        load()
        return True
    except Exception:  # noqa: BLE001
        return False


def lib():
    try:
        return load()
    except Exception as e:  # noqa: BLE001
        raise NativeError(
            'native library {} is not loadable ({}); build it with '
            '`python -m pytorch_distributed_mnist_b200.ops._build`'.format(lib_path(), e))


def check(status, what=''):
    if status != 0:
        msg = lib().pdm_last_error().decode('utf-8', 'replace')
        raise NativeError('{} failed with status {}: {}'.format(what or 'native call', status, msg))


def ptr(t):
    """Device/host pointer of a tensor as a c_void_p-compatible int (None -> NULL)."""
    if t is None:
        return None
    return t.data_ptr()


def stream_ptr(stream=None):
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream


def dtype_code(t):
    import torch
    if t.dtype == torch.uint8:
        return PDM_U8
    if t.dtype == torch.float32:
        return PDM_F32
    if t.dtype == torch.int64:
        return PDM_I64
    if t.dtype == torch.int32:
        return PDM_I32
    raise TypeError('unsupported dtype {} for the native kernels'.format(t.dtype))
