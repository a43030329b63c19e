"""``NativeReducer`` -- Python face of the native gradient reducer (``csrc/reducer.cu``).

The role of c10d's ``Reducer`` for arbitrary models: autograd's post-accumulate
hook forwards the parameter index; the native side counts bucket readiness, keeps buckets in
index order, orders the communication stream after the compute stream and launches the fused
all-reduce + 1/W + optimizer kernel of every bucket that became launchable -- the peer-memory
kernel or the in-switch (NVLS) one.  Without a native communicator (gloo, CPU) it does the
bookkeeping only and reports which buckets to reduce with ``torch.distributed``.
"""
from __future__ import annotations

import ctypes

from ..ops import _lib

c_void_p = ctypes.c_void_p
PDM_ERR_STATE = -3


class ReducerConfig(ctypes.Structure):
    _fields_ = [('n_params', ctypes.c_int32), ('n_buckets', ctypes.c_int32),
                ('bucket_of', ctypes.POINTER(ctypes.c_int32)),
                ('bucket_lo', ctypes.POINTER(ctypes.c_int64)), ('bucket_hi', ctypes.POINTER(ctypes.c_int64)),
                ('grads_off', ctypes.c_size_t), ('params_off', ctypes.c_size_t),
                ('m_off', ctypes.c_size_t), ('v_off', ctypes.c_size_t),
                ('comm', c_void_p), ('nvls', c_void_p), ('world', ctypes.c_int32), ('device', ctypes.c_int32)]


@_lib.optional_decl('pdm_reducer_create')
def _d1(fn):
    fn.argtypes = [ctypes.POINTER(c_void_p), ctypes.POINTER(ReducerConfig)]


@_lib.optional_decl('pdm_reducer_arm')
def _d2(fn):
    fn.argtypes = [c_void_p]


@_lib.optional_decl('pdm_reducer_mark_ready')
def _d3(fn):
    fn.argtypes = [c_void_p, ctypes.c_int32, c_void_p, c_void_p, c_void_p,
                   ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int32)]


@_lib.optional_decl('pdm_reducer_state')
def _d4(fn):
    fn.argtypes = [c_void_p, ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int64), c_void_p, c_void_p,
                   ctypes.POINTER(ctypes.c_int32)]


@_lib.optional_decl('pdm_reducer_unmarked')
def _d6(fn):
    fn.argtypes = [c_void_p, c_void_p, ctypes.POINTER(ctypes.c_int32)]


@_lib.optional_decl('pdm_reducer_finalize')
def _d7(fn):
    fn.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int32)]


@_lib.optional_decl('pdm_reducer_ready_order')
def _d8(fn):
    fn.argtypes = [c_void_p, c_void_p, ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int32)]


@_lib.optional_decl('pdm_reducer_destroy')
def _d5(fn):
    fn.argtypes = [c_void_p]


class NativeReducer(object):

    def __init__(self, bucket_of, ranges, world=1, comm_handle=None, nvls_handle=None, offsets=None, device=-1):
        """``bucket_of[i]``: bucket of parameter ``i``; ``ranges[b] = (lo, hi)``: float range of bucket
        ``b`` in the flat buffers; ``offsets``: byte offsets (grads, params, exp_avg, exp_avg_sq) of the
        flat buffers in the communicator's data area (needed with a native communicator only)."""
        self._lib = _lib.lib()
        self.n_params, self.n_buckets = len(bucket_of), len(ranges)
        bo = (ctypes.c_int32 * self.n_params)(*[int(b) for b in bucket_of])
        lo = (ctypes.c_int64 * self.n_buckets)(*[int(r[0]) for r in ranges])
        hi = (ctypes.c_int64 * self.n_buckets)(*[int(r[1]) for r in ranges])
        g, p, m, v = offsets if offsets is not None else (0, 0, 0, 0)
        cfg = ReducerConfig(self.n_params, self.n_buckets, bo, lo, hi, int(g), int(p), int(m), int(v),
                            comm_handle, nvls_handle, int(world), int(device))
        self._h = c_void_p()
        _lib.check(self._lib.pdm_reducer_create(ctypes.byref(self._h), ctypes.byref(cfg)), 'pdm_reducer_create')
        self.native_comm = comm_handle is not None or nvls_handle is not None
        self._first, self._count = ctypes.c_int32(0), ctypes.c_int32(0)

    def arm(self):
        """Start of a training forward; False when the previous backward left buckets unreduced."""
        rc = self._lib.pdm_reducer_arm(self._h)
        if rc == PDM_ERR_STATE:
            return False
        _lib.check(rc, 'pdm_reducer_arm')
        return True

    def mark_ready(self, index, hyper=None, compute_stream=0, comm_stream=0):
        """Returns ``(first, count)``: the buckets launched (or, without a native communicator, to be
        reduced by the caller) because of this gradient."""
        _lib.check(self._lib.pdm_reducer_mark_ready(self._h, int(index), ctypes.byref(hyper) if hyper is not None else None,
                                                    compute_stream, comm_stream, ctypes.byref(self._first),
                                                    ctypes.byref(self._count)), 'pdm_reducer_mark_ready')
        return self._first.value, self._count.value

    def unmarked(self):
        """Indices of the parameters that received no gradient in this pass so far."""
        out = (ctypes.c_int32 * self.n_params)()
        n = ctypes.c_int32(0)
        _lib.check(self._lib.pdm_reducer_unmarked(self._h, out, ctypes.byref(n)), 'pdm_reducer_unmarked')
        return list(out)[:n.value]

    def finalize(self, hyper=None, compute_stream=0, comm_stream=0):
        """End of a backward pass (``find_unused_parameters``): count the unmarked parameters as ready and
        launch the remaining buckets.  Returns ``(first, count)`` like ``mark_ready``."""
        _lib.check(self._lib.pdm_reducer_finalize(self._h, ctypes.byref(hyper) if hyper is not None else None,
                                                  compute_stream, comm_stream, ctypes.byref(self._first),
                                                  ctypes.byref(self._count)), 'pdm_reducer_finalize')
        return self._first.value, self._count.value

    def ready_order(self):
        """``(order, held_back)``: parameter indices in the order they became ready in this pass, and whether
        a completed bucket had to wait for an earlier one."""
        out = (ctypes.c_int32 * self.n_params)()
        n, held = ctypes.c_int32(0), ctypes.c_int32(0)
        _lib.check(self._lib.pdm_reducer_ready_order(self._h, out, ctypes.byref(n), ctypes.byref(held)),
                   'pdm_reducer_ready_order')
        return list(out)[:n.value], bool(held.value)

    def state(self):
        nxt, launches, n_order = ctypes.c_int32(0), ctypes.c_int64(0), ctypes.c_int32(0)
        pending = (ctypes.c_int32 * self.n_buckets)()
        order = (ctypes.c_int32 * self.n_buckets)()
        _lib.check(self._lib.pdm_reducer_state(self._h, ctypes.byref(nxt), ctypes.byref(launches), pending, order,
                                               ctypes.byref(n_order)), 'pdm_reducer_state')
        return {'next': nxt.value, 'launches': launches.value, 'pending': list(pending),
                'order': list(order)[:n_order.value]}

    def close(self):
        if self._h:
            self._lib.pdm_reducer_destroy(self._h)
            self._h = c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass
