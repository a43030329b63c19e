"""Flat parameter / gradient / optimizer-state storage.

The reference keeps every tensor separate: DDP copies each gradient into a flat
bucket, all-reduces, copies back (4 launches), then Adam walks
the parameter list again (8 launches, K11).  Here a model owns ONE fp32 buffer
laid out as ``[params | grads | exp_avg | exp_avg_sq]``; every ``nn.Parameter``,
its ``.grad`` and its optimizer state are views into it.  Consequences:

* a gradient bucket is a contiguous slice -- no gather/scatter copies;
* the optimizer is one kernel over a slice, fused with the all-reduce;
* when the buffer lives in the communicator's symmetric region, peers read
  gradients and write parameters in place over NVLink (``parallel.peer``).

Layout rule: parameters in ``module.parameters()`` order, each starting on a
4-float (16-byte) boundary; for ``Net`` this gives exactly the
``PDM_NET_FLAT`` layout of ``include/pdmnist.h`` (weight at 0, bias at 7840).
``relayout`` moves them into another order (the reducer's bucket rebuild); the
native engines freeze the layout they were built on.

Sharded optimizer state.  The two-shot form of the fused all-reduce + optimizer
(``csrc/comm.cu``, buckets >= 256 KB) lets rank r update only slice r of a
bucket: parameters are written to every replica, but ``exp_avg`` /
``exp_avg_sq`` of a slice stay current only on the rank that owns it.  That is
invisible while training (a rank only ever reads its own slice) but a
checkpoint written by rank 0 (ref:249-254) needs all of it, so whoever launches
such an update calls ``mark_sharded`` and ``consolidate`` gathers the slices
before the state is read (``runner.run`` does this before every save).
"""
from __future__ import annotations

import torch

# pdm_comm_allreduce_optim (algo = 0) switches to the two-shot kernel at this bucket size
TWO_SHOT_MIN_BYTES = 256 * 1024


def _round_up(x, m):
    return (x + m - 1) // m * m


class FlatState(object):
    SECTIONS = ('params', 'grads', 'exp_avg', 'exp_avg_sq')

    def __init__(self, params, device=None, comm=None):
        params = [p for p in params]
        if not params:
            raise ValueError('FlatState needs at least one parameter')
        for p in params:
            if p.dtype != torch.float32:
                raise TypeError('FlatState supports fp32 parameters only (got {})'.format(p.dtype))
        self.param_list = params
        self.device = torch.device(device) if device is not None else params[0].device
        self.offsets = []
        off = 0
        for p in params:
            self.offsets.append(off)
            off = _round_up(off + p.numel(), 4)
        self.layout = list(range(len(params)))        # parameter indices in memory order (see relayout)
        self.layout_frozen = False                    # set by the native engines, which hold raw pointers into the layout
        self.numel = off                              # floats, multiple of 4
        self.section = _round_up(self.numel, 64)      # section stride (256-byte aligned)
        self.comm = comm
        total = self.section * len(self.SECTIONS)
        if comm is not None:
            self.storage = comm.data_tensor(total)    # symmetric memory, zeroed at creation
        else:
            self.storage = torch.zeros(total, dtype=torch.float32, device=self.device)
        for i, name in enumerate(self.SECTIONS):
            setattr(self, name, self.storage[i * self.section:i * self.section + self.numel])
        self.optimizer = None          # set by FusedOptimizer
        self.fused_applied = False     # a reducer already applied this step's update
        self.sharded = {}              # (lo, hi) -> world: buckets whose optimizer state is updated slice-wise
        self.state_stale = False       # a slice-wise update ran since the last consolidate()
        with torch.no_grad():
            for p, o in zip(params, self.offsets):
                view = self.params[o:o + p.numel()].view(p.shape)
                view.copy_(p.data)
                p.data = view
                p._pdm_flat = self
        self.attach_grads()

    # byte offsets of the sections inside the storage (for the native comm calls)
    def byte_offset(self, name):
        return self.SECTIONS.index(name) * self.section * 4

    def view_of(self, name, index):
        p, o = self.param_list[index], self.offsets[index]
        return getattr(self, name)[o:o + p.numel()].view(p.shape)

    def attach_grads(self):
        """Point every ``p.grad`` at its slice of the flat gradient buffer."""
        for i, p in enumerate(self.param_list):
            if p.requires_grad:
                p.grad = self.view_of('grads', i)

    def zero_grads(self):
        self.grads.zero_()

    def range_of(self, first, last):
        """Float range [lo, hi) covering the parameters at memory positions first..last (inclusive),
        4-aligned.  Positions index ``self.layout``; without a relayout they are the parameter indices."""
        a, b = self.layout[first], self.layout[last]
        lo = self.offsets[a]
        hi = _round_up(self.offsets[b] + self.param_list[b].numel(), 4)
        return lo, hi

    def relayout(self, layout):
        """Move the parameters (with gradients and optimizer state) into a new memory order: ``layout``
        lists the parameter indices front to back.  Used by the reducer's bucket rebuild: a
        bucket is a contiguous range, so regrouping parameters means moving them.  Every view handed out
        before (``p.data``, ``p.grad``, optimizer state) is re-created; call ``consolidate`` first when
        the optimizer state is sharded."""
        layout = [int(i) for i in layout]
        n = len(self.param_list)
        if sorted(layout) != list(range(n)):
            raise ValueError('relayout needs a permutation of 0..{}'.format(n - 1))
        if self.layout_frozen:
            raise RuntimeError('a native engine holds pointers into this layout; it cannot be changed')
        if self.state_stale:
            raise RuntimeError('consolidate() the sharded optimizer state before a relayout')
        new_off, off = [0] * n, 0
        for i in layout:
            new_off[i] = off
            off = _round_up(off + self.param_list[i].numel(), 4)
        assert off == self.numel
        with torch.no_grad():
            for name in self.SECTIONS:
                buf = getattr(self, name)
                old = buf.clone()
                buf.zero_()
                for i, p in enumerate(self.param_list):
                    k = p.numel()
                    buf[new_off[i]:new_off[i] + k].copy_(old[self.offsets[i]:self.offsets[i] + k])
            self.offsets = new_off
            self.layout = layout
            for i, p in enumerate(self.param_list):
                p.data = self.view_of('params', i)
        self.attach_grads()
        self.sharded = {}
        if self.optimizer is not None:
            self.optimizer._make_state()


    # -- sharded optimizer state (see the module docstring) ---------------------
    @staticmethod
    def slice_of(lo, hi, world, rank):
        """Float range of bucket [lo, hi) that ``rank`` owns in the two-shot kernel: the bucket's
        float4s are cut into ``world`` runs of ceil(n4 / world) (the last ones may be short or empty)."""
        n4 = (hi - lo) // 4
        per = (n4 + world - 1) // world
        a = min(n4, per * rank)
        b = min(n4, a + per)
        return lo + 4 * a, lo + 4 * b

    def mark_sharded(self, lo, hi, world, force=False):
        """Record that the fused update of floats [lo, hi) runs slice-wise over ``world`` ranks
        (``force``: the in-switch kernel slices every bucket, whatever its size)."""
        if world > 1 and (force or 4 * (hi - lo) >= TWO_SHOT_MIN_BYTES):
            self.sharded[(int(lo), int(hi))] = int(world)
            self.state_stale = True

    def consolidate(self, group=None):
        """Collective: make ``exp_avg`` / ``exp_avg_sq`` complete on every rank.  Call on all ranks
        (like ``ZeroRedundancyOptimizer.consolidate_state_dict``); a no-op when nothing is sharded."""
        import torch.distributed as dist
        if not self.sharded or not self.state_stale:
            return 0
        if not (dist.is_available() and dist.is_initialized()):
            self.state_stale = False
            return 0
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        via_host = self.exp_avg.is_cuda and dist.get_backend(group) != 'nccl'
        moved = 0
        with torch.no_grad():
            for (lo, hi), w in sorted(self.sharded.items()):
                if w != world:
                    raise RuntimeError('bucket [{}, {}) was sharded over {} ranks, the group has {}'.format(lo, hi, w, world))
                per = (((hi - lo) // 4) + world - 1) // world * 4
                mine_lo, mine_hi = self.slice_of(lo, hi, world, rank)
                for buf in (self.exp_avg, self.exp_avg_sq):
                    send = torch.zeros(per, dtype=buf.dtype, device='cpu' if via_host else buf.device)
                    send[:mine_hi - mine_lo].copy_(buf[mine_lo:mine_hi])
                    parts = [torch.empty_like(send) for _ in range(world)]
                    dist.all_gather(parts, send, group=group)
                    for r in range(world):
                        r_lo, r_hi = self.slice_of(lo, hi, world, r)
                        if r != rank and r_hi > r_lo:
                            buf[r_lo:r_hi].copy_(parts[r][:r_hi - r_lo])
                            moved += r_hi - r_lo
        self.state_stale = False
        return moved


def flat_of(params):
    """The FlatState shared by all ``params`` (in order), or None."""
    params = list(params)
    flats = {id(getattr(p, '_pdm_flat', None)) for p in params}
    if len(flats) != 1:
        return None
    flat = getattr(params[0], '_pdm_flat', None)
    if flat is None or len(flat.param_list) != len(params):
        return None
    if any(a is not b for a, b in zip(flat.param_list, params)):
        return None
    return flat
