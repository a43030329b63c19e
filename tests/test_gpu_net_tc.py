"""The persistent tensor-core training kernel (csrc/net_tc.cu, ops/net_tc.py) against plain PyTorch
evaluations of the same steps: fp64 for single steps, fp32 torch.optim trajectories
for runs; ragged batches (1, 24, 777), batches longer than one 512-row block (1300), both input modes
(resident pool / index plan, device memory / pinned host memory), bitwise determinism."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

MEAN, STD = 0.1307, 0.3081


def _lib():
    from pytorch_distributed_mnist_b200.ops import _lib, net_tc
    if not net_tc.supported(torch.device('cuda', 0)):
        pytest.skip('device cannot launch the 13-CTA cluster kernel')
    return _lib, net_tc


def _data(n, seed, sparse=True):
    g = torch.Generator().manual_seed(seed)
    x = torch.randint(0, 256, (n, 784), generator=g, dtype=torch.uint8)
    if sparse:
        x[torch.rand(n, 784, generator=g) < 0.6] = 0          # MNIST-like: mostly background
    y = torch.randint(0, 10, (n,), generator=g, dtype=torch.uint8)
    return x, y


def _params(seed, scale=0.05):
    g = torch.Generator().manual_seed(seed)
    p = torch.zeros(7852)
    p[:7850] = torch.randn(7850, generator=g) * scale
    return p


def _norm(x):
    return x.float().div(255).sub(MEAN).div(STD)


@pytest.mark.parametrize('batch', [1, 24, 64, 512, 777, 1300])
def test_single_step_matches_fp64(batch):
    _l, net_tc = _lib()
    dev = torch.device('cuda', 0)
    x, y = _data(batch, batch)
    params = _params(7)
    pd = params.to(dev)
    grad = torch.zeros(7852, device=dev)
    logits = torch.zeros(batch, 10, device=dev)
    step_out = torch.zeros(2, device=dev)
    metrics = torch.zeros(4, dtype=torch.float64, device=dev)
    h = _l.Hyper(opt=_l.PDM_OPT_NONE, lr=0.0, beta1=0.0, beta2=0.0, eps=0.0, weight_decay=0.0, step=1)
    net_tc.tc_run(x.to(dev), y.to(dev), pd, None, None, h, batch, nsteps=1, grad_out=grad, logits_out=logits,
                  step_out=step_out, metrics=metrics)
    torch.cuda.synchronize()
    w = params[:7840].view(10, 784).double().requires_grad_(True)
    b = params[7840:7850].double().requires_grad_(True)
    z = ((x.double() / 255 - MEAN) / STD) @ w.t() + b
    loss = F.cross_entropy(z, y.long())
    loss.backward()
    assert torch.equal(pd.cpu(), params)                       # PDM_OPT_NONE leaves the parameters alone
    # fp32-level agreement: the MMAs see exact products (pixels exact in bf16, fp32 operands as 3 bf16 terms)
    assert (logits.cpu().double() - z.detach()).abs().max().item() < 2e-5 * max(1.0, z.detach().abs().max().item())
    assert abs(step_out[0].item() - loss.item()) < 1e-5 * max(1.0, loss.item())
    assert int(step_out[1].item()) == int((z.argmax(1) == y.long()).sum())
    gw, gb = grad[:7840].view(10, 784).cpu().double(), grad[7840:7850].cpu().double()
    # softmax uses ex2.approx (<= 2e-6 relative on a probability): allow 1e-5 of the gradient scale
    assert (gw - w.grad).abs().max().item() < 1e-5 * w.grad.abs().max().item() + 1e-9
    assert (gb - b.grad).abs().max().item() < 1e-5 * b.grad.abs().max().item() + 1e-9
    m = metrics.cpu()
    assert int(m[2].item()) == batch and abs(m[0].item() - loss.item() * batch) < 1e-4 * max(1.0, loss.item() * batch)


def _torch_run(params, batches, opt):
    lin = torch.nn.Linear(784, 10).cuda()
    with torch.no_grad():
        lin.weight.copy_(params[:7840].view(10, 784))
        lin.bias.copy_(params[7840:7850])
    if opt == 'adam':
        o = torch.optim.Adam(lin.parameters(), lr=1e-3)
    else:
        o = torch.optim.SGD(lin.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4)
    losses, correct = [], []
    for x, y in batches:
        out = lin(_norm(x.cuda()))
        loss = F.cross_entropy(out, y.long().cuda())
        o.zero_grad()
        loss.backward()
        o.step()
        losses.append(loss.item())
        correct.append(int((out.argmax(1) == y.long().cuda()).sum()))
    return lin, losses, correct


@pytest.mark.parametrize('opt', ['adam', 'sgd'])
def test_pool_run_matches_torch_optimizer(opt):
    _l, net_tc = _lib()
    dev = torch.device('cuda', 0)
    B, nb, steps = 512, 6, 30
    x, y = _data(B * nb, 3)
    params = _params(11)
    pd, m, v = params.to(dev), torch.zeros(7852, device=dev), torch.zeros(7852, device=dev)
    step_out = torch.zeros(steps, 2, device=dev)
    if opt == 'adam':
        h = _l.Hyper(opt=_l.PDM_OPT_ADAM, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, step=1)
    else:
        h = _l.Hyper(opt=_l.PDM_OPT_SGD, lr=0.05, beta1=0.9, beta2=0.0, eps=0.0, weight_decay=1e-4, step=1)
    net_tc.tc_run(x.to(dev), y.to(dev), pd, m, v, h, B, nsteps=steps, step_out=step_out)
    torch.cuda.synchronize()
    batches = [(x[(i % nb) * B:(i % nb + 1) * B], y[(i % nb) * B:(i % nb + 1) * B]) for i in range(steps)]
    lin, losses, correct = _torch_run(params, batches, opt)
    so = step_out.cpu()
    torch.testing.assert_close(so[:, 0], torch.tensor(losses), rtol=2e-3, atol=2e-3)
    assert (so[:, 1] - torch.tensor(correct, dtype=torch.float32)).abs().max().item() <= 1
    torch.testing.assert_close(pd[:7840].view(10, 784), lin.weight.detach(), rtol=1e-3, atol=2e-4)
    torch.testing.assert_close(pd[7840:7850], lin.bias.detach(), rtol=1e-3, atol=2e-4)


def test_plan_mode_ragged_and_split_runs_are_bitwise_equal():
    """An index plan over a dataset (last step ragged); the same plan as one launch and as two launches
    (state written back and reloaded, Adam step index carried over) gives bit-identical parameters; and
    two identical runs are bit-identical (fixed summation orders everywhere)."""
    _l, net_tc = _lib()
    dev = torch.device('cuda', 0)
    n, B, count = 2000, 96, 1000
    x, y = _data(n, 5)
    idx = torch.randperm(n, generator=torch.Generator().manual_seed(9))[:count].to(torch.int32)
    params = _params(13)
    xd, yd, idxd = x.to(dev), y.to(dev), idx.to(dev)
    nsteps = (count + B - 1) // B

    def run(splits):
        pd, m, v = params.to(dev), torch.zeros(7852, device=dev), torch.zeros(7852, device=dev)
        so = torch.zeros(nsteps, 2, device=dev)
        done = 0
        for k in splits:
            h = _l.Hyper(opt=_l.PDM_OPT_ADAM, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, step=done + 1)
            lo = done * B
            hi = min(count, (done + k) * B)
            net_tc.tc_run(xd, yd, pd, m, v, h, B, row_index=idxd[lo:hi], count=hi - lo, step_out=so[done:])
            done += k
        torch.cuda.synchronize()
        return pd.cpu(), m.cpu(), v.cpu(), so.cpu()

    one = run([nsteps])
    again = run([nsteps])
    two = run([4, nsteps - 4])
    for a, b in zip(one, again):
        assert torch.equal(a, b), 'two identical runs differ'
    for a, b in zip(one, two):
        assert torch.equal(a, b), 'one launch and two launches differ'
    batches = [(x[idx[s:s + B].long()], y[idx[s:s + B].long()]) for s in range(0, count, B)]
    assert batches[-1][0].size(0) == count - (nsteps - 1) * B != B
    lin, losses, correct = _torch_run(params, batches, 'adam')
    torch.testing.assert_close(one[3][:, 0], torch.tensor(losses), rtol=2e-3, atol=2e-3)
    torch.testing.assert_close(one[0][:7840].view(10, 784), lin.weight.detach().cpu(), rtol=1e-3, atol=2e-4)


def test_rows_from_pinned_host_memory():
    """Zero-copy input: the kernel's producer warps read the rows straight from page-locked host memory."""
    _l, net_tc = _lib()
    dev = torch.device('cuda', 0)
    n, B, count = 4096, 512, 2048
    x, y = _data(n, 21)
    xp, yp = x.pin_memory(), y.pin_memory()
    idx = torch.randperm(n, generator=torch.Generator().manual_seed(2))[:count].to(torch.int32).to(dev)
    params = _params(17)
    h = _l.Hyper(opt=_l.PDM_OPT_ADAM, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, step=1)
    ring = torch.empty(_l.TC_RING_BYTES, dtype=torch.uint8, device=dev)
    res = []
    for src in ('host', 'host+feeder', 'device', 'device+feeder'):
        pd, m, v = params.to(dev), torch.zeros(7852, device=dev), torch.zeros(7852, device=dev)
        so = torch.zeros(count // B, 2, device=dev)
        kw = dict(row_index=idx, step_out=so, ring=ring if 'feeder' in src else None)
        if src.startswith('host'):
            net_tc.tc_run((xp.data_ptr(), n), (yp.data_ptr(), n), pd, m, v, h, B, **kw)
        else:
            net_tc.tc_run(x.to(dev), y.to(dev), pd, m, v, h, B, **kw)
        torch.cuda.synchronize()
        res.append((pd.cpu(), so.cpu()))
    # the feeder cluster (whole rows gathered into a device ring by a second cluster of the same launch)
    # changes how bytes travel, not what is computed: all four runs are bit-identical
    for r in res[1:]:
        assert torch.equal(res[0][0], r[0]) and torch.equal(res[0][1], r[1])
    assert not torch.equal(res[0][0], params)


def test_feeder_ring_with_ragged_and_long_batches():
    """Feeder mode over more blocks than ring slots, with a batch longer than one 512-row block and a ragged tail."""
    _l, net_tc = _lib()
    dev = torch.device('cuda', 0)
    n, B, count = 6000, 700, 5000
    x, y = _data(n, 41)
    xp, yp = x.pin_memory(), y.pin_memory()
    idx = torch.randperm(n, generator=torch.Generator().manual_seed(4))[:count].to(torch.int32).to(dev)
    params = _params(19)
    h = _l.Hyper(opt=_l.PDM_OPT_ADAM, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, step=1)
    ring = torch.empty(_l.TC_RING_BYTES, dtype=torch.uint8, device=dev)
    nsteps = (count + B - 1) // B
    res = []
    for use_ring in (False, True):
        pd, m, v = params.to(dev), torch.zeros(7852, device=dev), torch.zeros(7852, device=dev)
        so = torch.zeros(nsteps, 2, device=dev)
        net_tc.tc_run((xp.data_ptr(), n), (yp.data_ptr(), n), pd, m, v, h, B, row_index=idx, step_out=so,
                      ring=ring if use_ring else None)
        torch.cuda.synchronize()
        res.append((pd.cpu(), so.cpu()))
    assert torch.equal(res[0][0], res[1][0]) and torch.equal(res[0][1], res[1][1])


def test_engine_uses_the_persistent_kernel():
    """NetEngine hands resident-pool and gather passes to the cluster kernel as ONE launch."""
    _l, net_tc = _lib()
    import pytorch_distributed_mnist_b200 as pdm
    from pytorch_distributed_mnist_b200.ops.engine import NetEngine
    dev = torch.device('cuda', 0)
    torch.manual_seed(0)
    net = pdm.Net().to(dev)
    opt = pdm.FusedAdam(net.parameters(), lr=1e-3)
    eng = NetEngine(opt.flat, opt, max_batch=256, slots=4, workers=2)
    x, y = _data(1024, 33)
    ref = torch.nn.Linear(784, 10).to(dev)
    ref.load_state_dict(net.fc.state_dict())
    ropt = torch.optim.Adam(ref.parameters(), lr=1e-3)
    before = eng.launches
    eng.train_device(x.to(dev), y.to(dev), 256, 8)
    assert eng.launches - before == 1
    for i in range(8):
        s = (i % 4) * 256
        loss = F.cross_entropy(ref(_norm(x[s:s + 256].to(dev))), y[s:s + 256].long().to(dev))
        ropt.zero_grad()
        loss.backward()
        ropt.step()
    eng.set_dataset(x, y)
    eng.place_dataset('zerocopy')
    idx = torch.randperm(1024, generator=torch.Generator().manual_seed(1))
    before = eng.launches
    steps, ms = eng.train(idx, 256)
    assert eng.launches - before == 1 and steps.shape == (4, 2)
    losses = []
    for s in range(0, 1024, 256):
        sel = idx[s:s + 256]
        loss = F.cross_entropy(ref(_norm(x[sel].to(dev))), y[sel].long().to(dev))
        ropt.zero_grad()
        loss.backward()
        ropt.step()
        losses.append(loss.item())
    torch.testing.assert_close(torch.from_numpy(steps[:, 0].copy()), torch.tensor(losses), rtol=2e-3, atol=2e-3)
    torch.testing.assert_close(net.fc.weight.detach(), ref.weight.detach(), rtol=1e-3, atol=2e-4)
    assert opt.step_count == 12
    eng.close()
