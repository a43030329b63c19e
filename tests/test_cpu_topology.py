"""Interconnect discovery (parallel/topology.py): probes are optional, answers are conservative."""
import types

import pytest

from pytorch_distributed_mnist_b200.parallel import topology as T


class FakeNvml(object):
    """An 8-GPU NVSwitch box (18 links per GPU, all ending in switches) or a 2-GPU NVLink bridge."""
    NVML_NVLINK_DEVICE_TYPE_SWITCH = 2
    NVML_P2P_CAPS_INDEX_ATOMICS = 3
    NVML_P2P_STATUS_OK = 0

    def __init__(self, n, switch=True, links=18, dead_links=()):
        self.n, self.switch, self.links, self.dead = n, switch, links, set(dead_links)
        self.inited = 0

    def nvmlInit(self):
        self.inited += 1

    def nvmlShutdown(self):
        self.inited -= 1

    def nvmlDeviceGetHandleByPciBusId(self, bus):
        raise RuntimeError('not found')          # force the index fallback

    def nvmlDeviceGetHandleByIndex(self, i):
        return i

    def nvmlDeviceGetPciInfo(self, h):
        return types.SimpleNamespace(busId='00000000:{:02X}:00.0'.format(0x10 + h).encode())

    def nvmlDeviceGetNvLinkState(self, h, link):
        if link >= self.links:
            raise RuntimeError('NVML_ERROR_NOT_SUPPORTED')
        return 0 if (h, link) in self.dead else 1

    def nvmlDeviceGetNvLinkRemoteDeviceType(self, h, link):
        return 2 if self.switch else 0

    def nvmlDeviceGetNvLinkRemotePciInfo(self, h, link):
        peer = (h + 1 + link % (self.n - 1)) % self.n
        return types.SimpleNamespace(busId='00000000:{:02x}:00.0'.format(0x10 + peer))

    def nvmlDeviceGetP2PStatus(self, a, b, idx):
        return 0


def _runtime(n, blocked=()):
    def probe(topo):
        topo.__init__(n)
        for a in range(n):
            topo.names[a] = 'FAKE B200'
            for b in range(n):
                if a != b:
                    topo.access[a][b] = (a, b) not in blocked
        return True
    return probe


def test_no_devices_on_this_box():
    topo = T.discover(refresh=True)
    # the build box has no GPU: discovery must say so quietly
    import torch
    if not torch.cuda.is_available():
        assert topo.n == 0 and topo.kind() == 'none' and not topo.all_peer_access([0, 1])
        assert T.main() == 0


def test_nvswitch_box():
    nv = FakeNvml(8, switch=True)
    topo = T.discover(probes=(('cuda-runtime', _runtime(8)), ('nvml', lambda t: T._probe_nvml(t, nv))))
    assert topo.sources == ['cuda-runtime', 'nvml'] and nv.inited == 0
    assert topo.all_peer_access() and topo.kind() == 'nvswitch' and topo.kind([0, 3]) == 'nvswitch'
    assert topo.switch_links == [18] * 8 and topo.nvlinks[0][1] == 0 and topo.atomics[2][5] is True
    assert topo.plan_allreduce(31400) == 'one-shot' and topo.plan_allreduce(64 << 20) == 'two-shot'
    assert topo.plan_allreduce(64 << 20, devices=[0, 1]) == 'one-shot'      # two ranks: nothing to scatter
    text = topo.summary()
    assert text.startswith('interconnect: nvswitch (8 devices') and 'NVS18' in text and 'FAKE B200' in text


def test_direct_nvlink_pair_and_dead_links():
    nv = FakeNvml(2, switch=False, links=4, dead_links=[(0, 3)])
    topo = T.discover(probes=(('cuda-runtime', _runtime(2)), ('nvml', lambda t: T._probe_nvml(t, nv))))
    assert topo.kind() == 'nvlink' and topo.nvlinks[0][1] == 3 and topo.nvlinks[1][0] == 4
    assert 'NV3' in topo.summary()


def test_blocked_pair_means_no_peer_path():
    topo = T.discover(probes=(('cuda-runtime', _runtime(4, blocked=[(1, 2)])),))
    assert not topo.all_peer_access() and topo.kind() == 'none'
    assert topo.all_peer_access([0, 3]) and topo.kind([0, 3]) == 'pcie'       # no NVLink info: assume the slow path
    assert topo.plan_allreduce(64 * 1024, devices=[0, 1, 3]) == 'two-shot'
    assert not topo.all_peer_access([0, 9])


def test_failing_probe_is_skipped():
    def boom(topo):
        raise OSError('libnvidia-ml.so.1: cannot open shared object file')
    topo = T.discover(probes=(('cuda-runtime', _runtime(2)), ('nvml', boom)))
    assert topo.sources == ['cuda-runtime'] and topo.kind() == 'pcie' and topo.switch_links == [None, None]


def test_cli_flag():
    from pytorch_distributed_mnist_b200.cli import build_parser
    assert build_parser().parse_args(['--print-topology']).print_topology is True
    assert build_parser().parse_args([]).print_topology is False


def test_peer_comm_supported_uses_discovery(monkeypatch):
    import torch
    from pytorch_distributed_mnist_b200.ops import _lib
    from pytorch_distributed_mnist_b200.parallel import peer
    topo = T.discover(probes=(('cuda-runtime', _runtime(4, blocked=[(1, 2)])),))
    topo.sources = ['cuda-runtime']
    monkeypatch.setattr(torch.cuda, 'is_available', lambda: True)
    monkeypatch.setattr(torch.cuda, 'device_count', lambda: 4)
    monkeypatch.setattr(_lib, 'available', lambda: True)
    monkeypatch.setattr(T, 'discover', lambda *a, **k: topo)
    assert not peer.PeerComm.supported(4)                    # GPUs 1 and 2 cannot map each other
    assert peer.PeerComm.supported(2, devices=[0, 3])
    assert not peer.PeerComm.supported(8)                    # more ranks than devices
    # discovery unavailable -> falls back to asking the runtime pair by pair
    empty = T.Topology(0)
    monkeypatch.setattr(T, 'discover', lambda *a, **k: empty)
    asked = []
    monkeypatch.setattr(torch.cuda, 'can_device_access_peer', lambda a, b: asked.append((a, b)) or True)
    assert peer.PeerComm.supported(2) and (0, 1) in asked
