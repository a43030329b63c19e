"""Numerics of the fused Net kernels vs plain PyTorch fp32."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

MEAN, STD = 0.1307, 0.3081


def _norm(x_u8):
    return x_u8.float().div(255.0).sub(MEAN).div(STD)


def _flat_from(linear):
    flat = torch.zeros(7852, device='cuda')
    flat[:7840] = linear.weight.detach().reshape(-1)
    flat[7840:7850] = linear.bias.detach()
    return flat


def _data(batch, seed=0, dtype='u8'):
    g = torch.Generator().manual_seed(seed)
    x = torch.randint(0, 256, (batch, 784), generator=g, dtype=torch.uint8)
    y = torch.randint(0, 10, (batch,), generator=g, dtype=torch.int64)
    xd = x.cuda()
    if dtype == 'f32':
        xd = _norm(xd)
    return xd, y.cuda()


@pytest.fixture(scope='module')
def stepper():
    from pytorch_distributed_mnist_b200.ops.fused_step import NetStepper
    return NetStepper('cuda:0')


@pytest.mark.parametrize('batch', [1, 7, 24, 64, 512, 1000, 4099])
@pytest.mark.parametrize('dtype', ['u8', 'f32'])
def test_forward_matches_torch(batch, dtype):
    from pytorch_distributed_mnist_b200.ops import _lib
    torch.manual_seed(1)
    lin = torch.nn.Linear(784, 10).cuda()
    x, _ = _data(batch, dtype=dtype)
    out = torch.empty(batch, 10, device='cuda')
    _lib.check(_lib.lib().pdm_net_forward(x.data_ptr(), _lib.dtype_code(x), batch, _flat_from(lin).data_ptr(),
                                          out.data_ptr(), _lib.stream_ptr()))
    ref = F.linear(x if dtype == 'f32' else _norm(x), lin.weight, lin.bias)
    torch.testing.assert_close(out, ref, rtol=1e-4, atol=1e-4)


def test_u8_normalisation_is_bit_exact():
    """The in-kernel LUT must reproduce ToTensor+Normalize bit for bit: with W = one-hot rows
    the logits are single normalised pixels."""
    from pytorch_distributed_mnist_b200.ops import _lib
    x = torch.arange(256, dtype=torch.uint8).repeat(4)[:784].view(1, 784).cuda()
    flat = torch.zeros(7852, device='cuda')
    cols = [0, 100, 255, 300, 511, 600, 700, 783, 5, 17]
    for c, k in enumerate(cols):
        flat[c * 784 + k] = 1.0
    out = torch.empty(1, 10, device='cuda')
    _lib.check(_lib.lib().pdm_net_forward(x.data_ptr(), _lib.PDM_U8, 1, flat.data_ptr(), out.data_ptr(),
                                          _lib.stream_ptr()))
    # oracle on the CPU: torchvision's ToTensor/Normalize run there (true division; the CUDA
    # div-by-scalar kernel multiplies by a reciprocal and differs in the last bit)
    ref = _norm(x.cpu())[0, cols]
    assert torch.equal(out[0].cpu(), ref)


@pytest.mark.parametrize('batch', [1, 24, 64, 512, 777, 5000])
def test_eval_loss_and_accuracy(stepper, batch):
    torch.manual_seed(2)
    lin = torch.nn.Linear(784, 10).cuda()
    x, y = _data(batch, seed=batch)
    stepper.read_metrics(reset=True)
    logits = torch.empty(batch, 10, device='cuda')
    stepper.eval_step(x, y.to(torch.uint8), _flat_from(lin), logits_out=logits)
    loss_sum, correct, samples = stepper.read_metrics()
    ref_logits = F.linear(_norm(x), lin.weight, lin.bias)
    ref_loss = F.cross_entropy(ref_logits.double(), y, reduction='sum').item()
    assert samples == batch
    assert correct == int((ref_logits.argmax(1) == y).sum().item())
    assert abs(loss_sum - ref_loss) <= 1e-4 * max(1.0, abs(ref_loss))
    torch.testing.assert_close(logits, ref_logits, rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize('batch,ctas', [(1, 0), (24, 0), (64, 0), (64, 3), (512, 0), (512, 148), (512, 1), (1000, 0), (3000, 0)])
@pytest.mark.parametrize('dtype', ['u8', 'f32'])
def test_gradient_matches_autograd(stepper, batch, ctas, dtype):
    from pytorch_distributed_mnist_b200.ops.fused_step import no_opt_hyper
    torch.manual_seed(3)
    lin = torch.nn.Linear(784, 10).cuda()
    x, y = _data(batch, seed=batch + 1, dtype=dtype)
    flat = _flat_from(lin)
    before = flat.clone()
    grad = torch.full((7852,), 123.0, device='cuda')
    out = stepper.train_step(x, y, flat, None, None, no_opt_hyper(), grad_out=grad, ctas=ctas)
    xf = x if dtype == 'f32' else _norm(x)
    loss = F.cross_entropy(F.linear(xf, lin.weight, lin.bias), y)
    loss.backward()
    torch.testing.assert_close(grad[:7840].view(10, 784), lin.weight.grad, rtol=2e-4, atol=2e-6)
    torch.testing.assert_close(grad[7840:7850], lin.bias.grad, rtol=2e-4, atol=2e-6)
    assert torch.equal(grad[7850:], torch.zeros(2, device='cuda'))
    assert torch.equal(flat, before)                       # PDM_OPT_NONE leaves the parameters alone
    got = out.tolist()
    assert abs(got[0] - loss.item()) < 1e-4 * max(1.0, abs(loss.item()))
    logits = F.linear(xf, lin.weight, lin.bias)
    assert int(round(got[1])) == int((logits.argmax(1) == y).sum().item())


def test_gradient_is_deterministic(stepper):
    from pytorch_distributed_mnist_b200.ops.fused_step import no_opt_hyper
    torch.manual_seed(4)
    lin = torch.nn.Linear(784, 10).cuda()
    x, y = _data(512, seed=9)
    flat = _flat_from(lin)
    g1 = torch.empty(7852, device='cuda')
    g2 = torch.empty(7852, device='cuda')
    stepper.train_step(x, y, flat, None, None, no_opt_hyper(), grad_out=g1)
    stepper.train_step(x, y, flat, None, None, no_opt_hyper(), grad_out=g2)
    assert torch.equal(g1, g2)


@pytest.mark.parametrize('opt', ['adam', 'adam_wd', 'sgd', 'sgd_momentum'])
def test_optimizer_trajectory_matches_torch(stepper, opt):
    from pytorch_distributed_mnist_b200.ops.fused_step import adam_hyper, sgd_hyper
    torch.manual_seed(5)
    lin = torch.nn.Linear(784, 10).cuda()
    flat = _flat_from(lin)
    m = torch.zeros_like(flat)
    v = torch.zeros_like(flat)
    if opt == 'adam':
        ref_opt = torch.optim.Adam(lin.parameters(), lr=1e-3)
        mk = lambda t: adam_hyper(lr=1e-3, step=t)
    elif opt == 'adam_wd':
        ref_opt = torch.optim.Adam(lin.parameters(), lr=2e-3, weight_decay=1e-2)
        mk = lambda t: adam_hyper(lr=2e-3, weight_decay=1e-2, step=t)
    elif opt == 'sgd':
        ref_opt = torch.optim.SGD(lin.parameters(), lr=1e-2)
        mk = lambda t: sgd_hyper(lr=1e-2, step=t)
    else:
        ref_opt = torch.optim.SGD(lin.parameters(), lr=1e-2, momentum=0.9, weight_decay=1e-4)
        mk = lambda t: sgd_hyper(lr=1e-2, momentum=0.9, weight_decay=1e-4, step=t)
    losses, ref_losses = [], []
    for t in range(1, 31):
        x, y = _data(64, seed=100 + t)
        out = stepper.train_step(x, y, flat, m, v, mk(t))
        losses.append(out[0].item())
        loss = F.cross_entropy(F.linear(_norm(x), lin.weight, lin.bias), y)
        ref_opt.zero_grad()
        loss.backward()
        ref_opt.step()
        ref_losses.append(loss.item())
    torch.testing.assert_close(torch.tensor(losses), torch.tensor(ref_losses), rtol=2e-3, atol=2e-3)
    torch.testing.assert_close(flat[:7840].view(10, 784), lin.weight.detach(), rtol=1e-3, atol=2e-4)
    torch.testing.assert_close(flat[7840:7850], lin.bias.detach(), rtol=1e-3, atol=2e-4)
    assert torch.equal(flat[7850:], torch.zeros(2, device='cuda'))


def test_metrics_accumulate_on_device(stepper):
    from pytorch_distributed_mnist_b200.ops.fused_step import adam_hyper
    torch.manual_seed(6)
    lin = torch.nn.Linear(784, 10).cuda()
    flat = _flat_from(lin)
    m, v = torch.zeros_like(flat), torch.zeros_like(flat)
    stepper.read_metrics(reset=True)
    total, n = 0.0, 0
    for t in range(1, 6):
        b = [64, 24, 512, 1, 100][t - 1]
        x, y = _data(b, seed=200 + t)
        out = stepper.train_step(x, y, flat, m, v, adam_hyper(step=t))
        total += out[0].item() * b
        n += b
    loss_sum, correct, samples = stepper.read_metrics()
    assert samples == n
    assert abs(loss_sum - total) < 1e-3 * max(1.0, total)
    assert 0 <= correct <= n


def test_bad_arguments_are_rejected(stepper):
    from pytorch_distributed_mnist_b200.ops import _lib
    from pytorch_distributed_mnist_b200.ops.fused_step import adam_hyper
    flat = torch.zeros(7852, device='cuda')
    x, y = _data(8)
    with pytest.raises(_lib.NativeError, match='Adam needs m and v'):
        stepper.train_step(x, y, flat, None, None, adam_hyper())
    with pytest.raises(_lib.NativeError, match='ctas'):
        stepper.train_step(x, y, flat, flat.clone(), flat.clone(), adam_hyper(), ctas=1000)
