"""Python face of the native peer-memory communicator (``csrc/comm.cu``).

Bootstrap: the ``tcp://`` rendezvous contract of the reference
(ref:167-168, ref:323-327) is kept -- ``dist.init_process_group`` still builds
the TCP store -- but the store is used only to exchange 64-byte CUDA IPC
handles.  After that every rank has every peer's symmetric region mapped and
all gradient traffic is SM loads/stores over NVLink issued by our kernels; no
NCCL call is made on the training path.

``PeerComm.supported()`` is the topology probe: every pair of visible devices
in the job must report P2P access (on an HGX B200 box that is NVLink through
NVSwitch for all 8 GPUs).  When it is False the caller falls back to
``GroupComm`` (stock ``torch.distributed`` collectives).
"""
from __future__ import annotations

import ctypes
import itertools
import time

import torch
import torch.distributed as dist

from ..ops import _lib

_counter = itertools.count()


class _CudaBlock(object):
    """Minimal ``__cuda_array_interface__`` holder so torch can view raw device memory."""

    def __init__(self, ptr, nbytes, owner):
        self.owner = owner
        self.__cuda_array_interface__ = {
            'shape': (nbytes,), 'typestr': '|u1', 'data': (int(ptr), False), 'version': 3, 'strides': None,
        }


def default_store():
    from torch.distributed import distributed_c10d
    return distributed_c10d._get_default_store()


class PeerComm(object):

    def __init__(self, rank, world, device, data_bytes, store=None, timeout_s=30.0, tag=None):
        self.rank, self.world = int(rank), int(world)
        self.device = torch.device(device)
        self._lib = _lib.lib()
        self._h = ctypes.c_void_p()
        _lib.check(self._lib.pdm_comm_create(ctypes.byref(self._h), self.rank, self.world,
                                             self.device.index or 0, int(data_bytes)), 'pdm_comm_create')
        self._lib.pdm_comm_set_timeout(self._h, float(timeout_s))
        self.data_bytes = int(self._lib.pdm_comm_data_bytes(self._h))
        self._data_ptr = int(self._lib.pdm_comm_data(self._h))
        self._cursor = 0
        self._block = _CudaBlock(self._data_ptr, max(self.data_bytes, 1), self)
        self._bytes = torch.as_tensor(self._block, device=self.device) if self.data_bytes else None
        if self.world > 1:
            self._exchange(store if store is not None else default_store(), tag)

    # -- bootstrap ----------------------------------------------------------
    def _exchange(self, store, tag):
        tag = tag if tag is not None else 'pdm/comm{}'.format(next(_counter))
        mine = ctypes.create_string_buffer(64)
        _lib.check(self._lib.pdm_comm_handle(self._h, mine), 'pdm_comm_handle')
        store.set('{}/handle/{}'.format(tag, self.rank), mine.raw)
        blob = b''.join(bytes(store.get('{}/handle/{}'.format(tag, r))) for r in range(self.world))
        if len(blob) != 64 * self.world:
            raise RuntimeError('IPC handle exchange returned {} bytes, expected {}'.format(len(blob), 64 * self.world))
        _lib.check(self._lib.pdm_comm_connect(self._h, blob), 'pdm_comm_connect')
        # nobody may tear down or reuse keys before everyone has mapped everyone
        store.add('{}/mapped'.format(tag), 1)
        while int(store.add('{}/mapped'.format(tag), 0)) < self.world:
            time.sleep(0.001)

    @staticmethod
    def supported(world, devices=None):
        """True when every device pair of the job can map each other's memory."""
        if not torch.cuda.is_available() or not _lib.available():
            return False
        n = torch.cuda.device_count()
        devices = list(devices) if devices is not None else list(range(world))
        if world > _lib.PDM_MAX_RANKS or any(d >= n for d in devices):
            return False
        from . import topology
        topo = topology.discover()
        if topo.n == n and 'cuda-runtime' in topo.sources:
            return topo.all_peer_access(devices)
        for a in devices:               # discovery unavailable: ask the runtime directly
            for b in devices:
                if a != b and not torch.cuda.can_device_access_peer(a, b):
                    return False
        return True

    # -- memory -------------------------------------------------------------
    def data_tensor(self, numel_f32):
        """Carve ``numel_f32`` floats (256-byte aligned) out of the symmetric data area."""
        nbytes = (int(numel_f32) * 4 + 255) // 256 * 256
        if self._cursor + nbytes > self.data_bytes:
            raise RuntimeError('symmetric region exhausted: need {} more bytes, {} left'
                               .format(nbytes, self.data_bytes - self._cursor))
        lo = self._cursor
        self._cursor += nbytes
        t = self._bytes[lo:lo + int(numel_f32) * 4].view(torch.float32)
        t._pdm_offset = lo
        return t

    def offset_of(self, tensor):
        """Byte offset of a tensor carved from this region."""
        off = tensor.data_ptr() - self._data_ptr
        if off < 0 or off >= self.data_bytes:
            raise ValueError('tensor is not inside this communicator\'s symmetric region')
        return off

    # -- collectives --------------------------------------------------------
    def barrier(self, stream=None):
        _lib.check(self._lib.pdm_comm_barrier(self._h, _lib.stream_ptr(stream)), 'pdm_comm_barrier')

    def broadcast(self, tensor, root=0, stream=None):
        nbytes = (tensor.numel() * tensor.element_size() + 15) // 16 * 16
        _lib.check(self._lib.pdm_comm_broadcast(self._h, self.offset_of(tensor), nbytes, int(root),
                                                _lib.stream_ptr(stream)), 'pdm_comm_broadcast')

    def allreduce_optim(self, grads, out, exp_avg, exp_avg_sq, hyper, grad_scale=None, algo=0,
                        max_ctas=0, stream=None):
        """Fused all-reduce + scale + optimizer on 4-float-aligned slices of the region."""
        n = grads.numel()
        scale = (1.0 / self.world) if grad_scale is None else grad_scale
        _lib.check(self._lib.pdm_comm_allreduce_optim(
            self._h, self.offset_of(grads), self.offset_of(out),
            self.offset_of(exp_avg) if exp_avg is not None else 0,
            self.offset_of(exp_avg_sq) if exp_avg_sq is not None else 0,
            n, ctypes.byref(hyper), scale, int(algo), int(max_ctas), _lib.stream_ptr(stream)),
            'pdm_comm_allreduce_optim')

    def allreduce(self, src, dst, average=True, algo=0, stream=None):
        """Plain all-reduce of ``src`` into ``dst`` (both inside the region)."""
        h = _lib.Hyper(opt=_lib.PDM_OPT_NONE, lr=0, beta1=0, beta2=0, eps=0, weight_decay=0, step=1)
        self.allreduce_optim(src, dst, None, None, h, (1.0 / self.world) if average else 1.0, algo, 0, stream)

    def peers_struct(self, advance=True):
        p = _lib.Peers()
        _lib.check(self._lib.pdm_comm_peers(self._h, ctypes.byref(p), 1 if advance else 0), 'pdm_comm_peers')
        return p

    def status(self):
        s = ctypes.c_int(0)
        _lib.check(self._lib.pdm_comm_status(self._h, ctypes.byref(s)), 'pdm_comm_status')
        return s.value

    def check_status(self):
        s = self.status()
        if s:
            raise RuntimeError('rank {}: peer rank {} did not reach a collective within the timeout '
                               '(dead or hung process?)'.format(self.rank, s - 1))

    @property
    def handle(self):
        return self._h

    def close(self):
        if self._h:
            if self.world > 1 and dist.is_available() and dist.is_initialized():
                try:
                    dist.barrier()   # peers may still be reading our region
                except Exception:  # noqa: BLE001
                    pass
            self._lib.pdm_comm_destroy(self._h)
            self._h = ctypes.c_void_p()


class GroupComm(object):
    """Stock ``torch.distributed`` fallback (gloo on CPU, nccl on GPU): the measured
    baseline path and the path of BASELINE config 1 (SURVEY 5.8 'Fallback')."""

    def __init__(self, group=None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0

    def barrier(self, stream=None):
        if self.world > 1:
            dist.barrier(self.group)

    def broadcast(self, tensor, root=0, stream=None):
        if self.world > 1:
            dist.broadcast(tensor, root, group=self.group)

    def allreduce_async(self, tensor):
        """Launch SUM all-reduce; returns a handle with ``wait()`` (None when world == 1)."""
        if self.world > 1:
            return dist.all_reduce(tensor, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        return None

    def check_status(self):
        pass

    def close(self):
        pass
