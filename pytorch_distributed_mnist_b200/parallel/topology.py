"""Interconnect discovery for the peer-memory communicator (SURVEY K16 / 5.8 "bootstrap").

The reference never looks at the machine: NCCL does that inside ``ncclCommInitRank`` (topology
search, ring/tree construction) the first time DDP issues a collective (ref:188-189).  Our
communicator has no rings to build -- behind an NVSwitch every peer is one hop away at full
bandwidth -- but it still has to know three things before it maps anybody's memory:

* can every pair of GPUs in the job load/store each other's memory (else: ``--comm group``),
* is the path NVLink (and through a switch) or PCIe, which decides whether the one-shot or the
  two-shot all-reduce wins for a given bucket size (``plan_allreduce``),
* do the devices support multicast objects (NVLS): the reducer then reduces large models in the
  switch (``parallel/nvls.py``, ``parallel/ddp.py``: >= 8 MB at >= 4 GPUs).

Sources: the CUDA runtime through torch (peer access), NVML (per-link state, remote device type,
P2P capability matrix) and the driver attribute for multicast.  Every source is optional: a probe
that is missing or fails leaves its fields ``None`` and never raises, so the module also works on
the GPU-less build box and under ``CUDA_VISIBLE_DEVICES`` remapping.

    python -m pytorch_distributed_mnist_b200.parallel.topology        # print what was found
"""
from __future__ import annotations

import json

NVLINK_MAX_LINKS = 18
# comm.cu switches from the one-shot to the two-shot kernel at 256 KB per bucket (algo = 0), which is
# where the two curves cross on the NVSwitch boxes this was measured on (profiles/README.md section 2).
# The other rows are estimates for machines we could not measure: fewer links or PCIe make every
# replicated byte more expensive, so the crossover comes earlier.
TWO_SHOT_MIN_BYTES = {'nvswitch': 256 * 1024, 'nvlink': 128 * 1024, 'pcie': 32 * 1024, 'none': 0}


class Topology(object):
    """Plain data: what the probes found for ``n`` visible devices."""

    def __init__(self, n=0):
        self.n = int(n)
        self.names = [None] * self.n
        self.pci = [None] * self.n
        self.access = [[(a == b) for b in range(self.n)] for a in range(self.n)]   # can a map b's memory
        self.nvlinks = [[None] * self.n for _ in range(self.n)]    # NVLink links a->b (None unknown)
        self.atomics = [[None] * self.n for _ in range(self.n)]    # native atomics over the path
        self.switch_links = [None] * self.n                        # links of device a that end in an NVSwitch
        self.multicast = [None] * self.n                           # CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED
        self.sources = []                                          # which probes answered

    # -- queries ------------------------------------------------------------
    def all_peer_access(self, devices=None):
        devices = list(range(self.n)) if devices is None else list(devices)
        if any(d < 0 or d >= self.n for d in devices):
            return False
        return all(self.access[a][b] for a in devices for b in devices if a != b)

    def kind(self, devices=None):
        """'nvswitch' | 'nvlink' | 'pcie' | 'none' for the given device set (worst pair decides)."""
        devices = list(range(self.n)) if devices is None else list(devices)
        if len(devices) < 2:
            return 'none'
        if not self.all_peer_access(devices):
            return 'none'
        if all((self.switch_links[d] or 0) > 0 for d in devices):
            return 'nvswitch'
        pairs = [(a, b) for a in devices for b in devices if a != b]
        if all((self.nvlinks[a][b] or 0) > 0 for a, b in pairs):
            return 'nvlink'
        return 'pcie'

    def plan_allreduce(self, nbytes, devices=None):
        """'one-shot' or 'two-shot' for a bucket of ``nbytes`` on this interconnect.  Advisory: the
        native kernel picks by size alone (``algo = 0`` in ``pdm_comm_allreduce_optim``), which is the
        same answer on an NVSwitch machine; callers can pass ``algo=1|2`` to follow this plan elsewhere."""
        world = self.n if devices is None else len(list(devices))
        if world <= 2:
            return 'one-shot'
        return 'two-shot' if nbytes >= TWO_SHOT_MIN_BYTES[self.kind(devices)] else 'one-shot'

    def as_dict(self):
        return {k: getattr(self, k) for k in ('n', 'names', 'pci', 'access', 'nvlinks', 'atomics', 'switch_links',
                                              'multicast', 'sources')}

    def summary(self, devices=None):
        devices = list(range(self.n)) if devices is None else list(devices)
        out = ['interconnect: {} ({} device{}; probes: {})'.format(
            self.kind(devices), len(devices), '' if len(devices) == 1 else 's', ', '.join(self.sources) or 'none')]
        if not devices:
            return out[0] + ' -- no CUDA devices: gradients go through torch.distributed (--comm group)'
        head = '      ' + ' '.join('GPU{:<3d}'.format(d) for d in devices)
        out.append(head)
        for a in devices:
            cells = []
            for b in devices:
                if a == b:
                    cells.append('  X   ')
                elif not self.access[a][b]:
                    cells.append('  -   ')
                elif self.nvlinks[a][b]:
                    cells.append('NV{:<4d}'.format(self.nvlinks[a][b]))
                elif self.switch_links[a]:
                    cells.append('NVS{:<3d}'.format(self.switch_links[a]))
                else:
                    cells.append(' P2P  ')
            out.append('GPU{:<2d} '.format(a) + ' '.join(cells))
        extra = []
        if any(v is not None for v in self.multicast):
            extra.append('multicast (NVLS) objects: {}'.format(
                'supported' if all(self.multicast[d] for d in devices) else 'not supported'))
        if self.names and self.names[0]:
            extra.append('device: {}'.format(self.names[devices[0]] if devices else self.names[0]))
        out.extend(extra)
        return '\n'.join(out)


# ---------------------------------------------------------------------------
# probes
# ---------------------------------------------------------------------------
def _probe_torch(topo):
    import torch
    if not torch.cuda.is_available():
        return False
    n = torch.cuda.device_count()
    if topo.n == 0:
        topo.__init__(n)
    for a in range(topo.n):
        props = torch.cuda.get_device_properties(a)
        topo.names[a] = props.name
        try:
            topo.pci[a] = '{:08x}:{:02x}:{:02x}.0'.format(props.pci_domain_id, props.pci_bus_id, props.pci_device_id)
        except AttributeError:
            pass
        for b in range(topo.n):
            if a != b:
                topo.access[a][b] = bool(torch.cuda.can_device_access_peer(a, b))
    return True


def _probe_nvml(topo, nvml=None):
    """Per-link NVLink state + P2P capability matrix.  ``nvml`` is injectable for tests."""
    if nvml is None:
        import pynvml as nvml
    nvml.nvmlInit()
    try:
        handles = []
        for a in range(topo.n):
            h = None
            if topo.pci[a]:
                try:
                    h = nvml.nvmlDeviceGetHandleByPciBusId(topo.pci[a].encode() if isinstance(topo.pci[a], str) else topo.pci[a])
                except Exception:  # noqa: BLE001 - fall back to the index
                    h = None
            if h is None:
                h = nvml.nvmlDeviceGetHandleByIndex(a)
            handles.append(h)
        bus_of = {}
        for a, h in enumerate(handles):
            try:
                bus = nvml.nvmlDeviceGetPciInfo(h).busId
                bus_of[(bus.decode() if isinstance(bus, bytes) else bus).lower()] = a
            except Exception:  # noqa: BLE001
                pass
        switch_type = getattr(nvml, 'NVML_NVLINK_DEVICE_TYPE_SWITCH', 2)
        for a, h in enumerate(handles):
            to_switch, per_peer = 0, {}
            for link in range(NVLINK_MAX_LINKS):
                try:
                    if not nvml.nvmlDeviceGetNvLinkState(h, link):
                        continue
                except Exception:  # noqa: BLE001 - link not present on this part
                    continue
                try:
                    if nvml.nvmlDeviceGetNvLinkRemoteDeviceType(h, link) == switch_type:
                        to_switch += 1
                        continue
                except Exception:  # noqa: BLE001
                    pass
                try:
                    bus = nvml.nvmlDeviceGetNvLinkRemotePciInfo(h, link).busId
                    b = bus_of.get((bus.decode() if isinstance(bus, bytes) else bus).lower())
                    if b is not None:
                        per_peer[b] = per_peer.get(b, 0) + 1
                except Exception:  # noqa: BLE001
                    pass
            topo.switch_links[a] = to_switch
            for b in range(topo.n):
                if a != b:
                    topo.nvlinks[a][b] = per_peer.get(b, 0)
        idx_atomics = getattr(nvml, 'NVML_P2P_CAPS_INDEX_ATOMICS', 3)
        ok = getattr(nvml, 'NVML_P2P_STATUS_OK', 0)
        for a in range(topo.n):
            for b in range(topo.n):
                if a == b:
                    continue
                try:
                    topo.atomics[a][b] = nvml.nvmlDeviceGetP2PStatus(handles[a], handles[b], idx_atomics) == ok
                except Exception:  # noqa: BLE001
                    pass
    finally:
        try:
            nvml.nvmlShutdown()
        except Exception:  # noqa: BLE001
            pass
    return True


def _probe_multicast(topo):
    from cuda.bindings import driver
    attr = driver.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED
    (err,) = driver.cuInit(0)
    if int(err) != 0:
        return False
    for a in range(topo.n):
        err, dev = driver.cuDeviceGet(a)
        if int(err) != 0:
            continue
        err, val = driver.cuDeviceGetAttribute(attr, dev)
        if int(err) == 0:
            topo.multicast[a] = bool(val)
    return True


_cache = None


def discover(refresh=False, probes=None):
    """Run the probes once per process and cache the result."""
    global _cache
    if _cache is not None and not refresh and probes is None:
        return _cache
    topo = Topology(0)
    for name, fn in (probes or (('cuda-runtime', _probe_torch), ('nvml', _probe_nvml), ('driver-multicast', _probe_multicast))):
        try:
            if fn(topo):
                topo.sources.append(name)
        except Exception:  # noqa: BLE001 - a missing probe must never stop a job
            continue
        if topo.n == 0:
            break          # no CUDA devices: nothing for the later probes to describe
    if probes is None:
        _cache = topo
    return topo


def main():
    topo = discover(refresh=True)
    if topo.n == 0:
        print('no CUDA devices visible: the peer-memory communicator is unavailable; --comm group '
              '(torch.distributed gloo/nccl) is used')
        return 0
    print(topo.summary())
    print(json.dumps({'kind': topo.kind(), 'all_peer_access': topo.all_peer_access(),
                      'allreduce_plan': {str(nb): topo.plan_allreduce(nb) for nb in (31400, 1 << 20, 64 << 20)}}))
    return 0


if __name__ == '__main__':
    raise SystemExit(main())
