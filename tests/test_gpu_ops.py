"""Flat optimizer kernel, fused cross-entropy, single-rank communicator."""
import ctypes

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('n', [4, 7850, 7852, 100003, 1 << 20])
@pytest.mark.parametrize('kind', ['adam', 'sgd'])
def test_flat_optimizer_matches_torch(n, kind):
    from pytorch_distributed_mnist_b200.ops import _lib
    torch.manual_seed(0)
    p = torch.randn(n, device='cuda')
    ref_p = p.clone().requires_grad_(True)
    m = torch.zeros(n, device='cuda')
    v = torch.zeros(n, device='cuda')
    if kind == 'adam':
        ropt = torch.optim.Adam([ref_p], lr=3e-3, weight_decay=1e-3)
    else:
        ropt = torch.optim.SGD([ref_p], lr=3e-2, momentum=0.9, weight_decay=1e-3)
    for t in range(1, 6):
        g = torch.randn(n, device='cuda')
        h = _lib.Hyper(opt=_lib.PDM_OPT_ADAM if kind == 'adam' else _lib.PDM_OPT_SGD, lr=3e-3 if kind == 'adam' else 3e-2,
                       beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=1e-3, step=t)
        _lib.check(_lib.lib().pdm_optim_step(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), n,
                                             ctypes.byref(h), 1.0, _lib.stream_ptr()))
        ref_p.grad = g.clone()
        ropt.step()
    torch.testing.assert_close(p, ref_p.detach(), rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize('batch,classes', [(1, 10), (64, 10), (512, 10), (3000, 10), (100, 32)])
@pytest.mark.parametrize('ydtype', [torch.int64, torch.uint8])
def test_fused_cross_entropy(batch, classes, ydtype):
    from pytorch_distributed_mnist_b200.ops.loss import cross_entropy
    torch.manual_seed(1)
    z = (torch.randn(batch, classes, device='cuda') * 3).requires_grad_(True)
    y = torch.randint(0, classes, (batch,), device='cuda')
    metrics = torch.zeros(4, dtype=torch.float64, device='cuda')
    loss = cross_entropy(z, y.to(ydtype), metrics)
    loss.backward()
    zr = z.detach().clone().requires_grad_(True)
    ref = F.cross_entropy(zr, y)
    ref.backward()
    torch.testing.assert_close(loss, ref, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(z.grad, zr.grad, rtol=1e-4, atol=1e-7)
    got = metrics.tolist()
    assert int(got[2]) == batch
    assert int(got[1]) == int((zr.argmax(1) == y).sum())
    assert abs(got[0] - ref.item() * batch) < 1e-3 * max(1.0, ref.item() * batch)


def test_single_rank_communicator_is_a_fused_optimizer():
    """world == 1: pdm_comm_allreduce_optim degenerates to the flat optimizer, the broadcast
    and the barrier to no-ops -- the same code path the multi-GPU job runs."""
    from pytorch_distributed_mnist_b200.ops import _lib
    from pytorch_distributed_mnist_b200.parallel.peer import PeerComm
    n = 4096
    comm = PeerComm(0, 1, torch.device('cuda', 0), data_bytes=n * 4 * 4 + 4096)
    p, g, m, v = (comm.data_tensor(n) for _ in range(4))
    torch.manual_seed(2)
    p.copy_(torch.randn(n))
    g.copy_(torch.randn(n))
    ref_p = p.clone().requires_grad_(True)
    ropt = torch.optim.Adam([ref_p], lr=1e-2)
    ref_p.grad = g.clone()
    ropt.step()
    comm.barrier()
    comm.broadcast(p)
    h = _lib.Hyper(opt=_lib.PDM_OPT_ADAM, lr=1e-2, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, step=1)
    comm.allreduce_optim(g, p, m, v, h)
    torch.cuda.synchronize()
    torch.testing.assert_close(p, ref_p.detach(), rtol=1e-5, atol=1e-6)
    assert comm.status() == 0
    comm.close()


def test_net_forward_module_uses_native_kernel():
    from pytorch_distributed_mnist_b200 import Net
    torch.manual_seed(3)
    net = Net().cuda()
    x = torch.randn(33, 1, 28, 28, device='cuda')
    out = net(x)
    ref = F.linear(x.view(33, -1), net.fc.weight, net.fc.bias)
    torch.testing.assert_close(out, ref, rtol=1e-4, atol=1e-4)
    out.sum().backward()
    assert net.fc.weight.grad is not None and net.fc.bias.grad is not None


def test_find_unused_parameters_on_the_peer_path():
    """The native reducer's end-of-backward pass on a GPU: the branch that received no gradient is counted
    as ready, every bucket kernel still runs (fused SGD inside), parameters follow torch.optim.SGD."""
    import pytorch_distributed_mnist_b200 as pdm

    class TwoBranch(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.a = torch.nn.Linear(600, 500)
            self.b = torch.nn.Linear(600, 500)
            self.c = torch.nn.Linear(500, 4)
            self.never = torch.nn.Linear(3, 3)

        def forward(self, x, use_b):
            return self.c(torch.relu(self.b(x) if use_b else self.a(x)))

    torch.manual_seed(5)
    ours = pdm.DistributedDataParallel(TwoBranch().cuda(), device_ids=[0], find_unused_parameters=True, bucket_cap_mb=1,
                                       rebuild_buckets=False)      # the regrouping has its own test (gloo, CPU)
    torch.manual_seed(5)
    ref = TwoBranch().cuda()
    assert ours.uses_peer and len(ours.buckets) >= 2
    o_opt = pdm.FusedSGD(ours.parameters(), lr=0.05)
    r_opt = torch.optim.SGD(ref.parameters(), lr=0.05)
    g = torch.Generator().manual_seed(11)
    for i in range(4):
        x = torch.randn(16, 600, generator=g).cuda()
        y = torch.randint(0, 4, (16,), generator=g).cuda()
        for m, o in ((ours, o_opt), (ref, r_opt)):
            o.zero_grad()
            F.cross_entropy(m(x, i % 2 == 1), y).backward()
            o.step()
    torch.cuda.synchronize()
    for (n, a), b in zip(ours.module.named_parameters(), ref.parameters()):
        torch.testing.assert_close(a, b, rtol=1e-4, atol=1e-5, msg=lambda m, n=n: n + ': ' + m)
    st = ours._reducer.state()
    assert ours._reducer.native_comm and st['next'] == len(ours.buckets)
    assert ours.stats['bucket_launches'] == 4 * len(ours.buckets)
    assert ours.stats['unused_parameters'] == 4
    ours.close()
