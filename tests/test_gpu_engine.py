"""Native epoch executor + fused trainer path on one GPU."""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

MEAN, STD = 0.1307, 0.3081


def _norm(x_u8):
    return x_u8.float().div(255.0).sub(MEAN).div(STD)


def _make(n=3000, seed=0):
    from pytorch_distributed_mnist_b200.data import mnist
    images, labels = mnist.synthetic_mnist(True, seed=seed, size=n)
    return torch.from_numpy(images.reshape(n, 784)), torch.from_numpy(labels)


def _build(lr=1e-3, opt='adam'):
    from pytorch_distributed_mnist_b200 import Net, FusedAdam, FusedSGD
    torch.manual_seed(0)
    net = Net().cuda()
    ref = torch.nn.Linear(784, 10).cuda()
    ref.load_state_dict(net.fc.state_dict())
    if opt == 'adam':
        return net, FusedAdam(net.parameters(), lr=lr), ref, torch.optim.Adam(ref.parameters(), lr=lr)
    return (net, FusedSGD(net.parameters(), lr=lr, momentum=0.9, weight_decay=1e-4), ref,
            torch.optim.SGD(ref.parameters(), lr=lr, momentum=0.9, weight_decay=1e-4))


@pytest.mark.parametrize('opt', ['adam', 'sgd'])
@pytest.mark.parametrize('batch,workers', [(64, 1), (100, 3)])
def test_engine_epoch_matches_torch_loop(opt, batch, workers):
    from pytorch_distributed_mnist_b200.ops.engine import NetEngine
    images, labels = _make(1000)
    net, fopt, ref, ropt = _build(opt=opt)
    eng = NetEngine(fopt.flat, fopt, max_batch=256, slots=4, workers=workers)
    eng.set_dataset(images, labels)
    idx = torch.randperm(1000, generator=torch.Generator().manual_seed(1))
    steps, ms = eng.train(idx, batch)
    assert steps.shape == ((1000 + batch - 1) // batch, 2)
    assert ms > 0
    ref_losses, ref_correct = [], []
    for s in range(0, 1000, batch):
        sel = idx[s:s + batch]
        x, y = _norm(images[sel].cuda()), labels[sel].long().cuda()
        out = ref(x)
        loss = F.cross_entropy(out, y)
        ropt.zero_grad()
        loss.backward()
        ropt.step()
        ref_losses.append(loss.item())
        ref_correct.append(int((out.argmax(1) == y).sum()))
    np.testing.assert_allclose(steps[:, 0], np.array(ref_losses), rtol=2e-3, atol=2e-3)
    assert np.abs(steps[:, 1] - np.array(ref_correct)).max() <= 1
    torch.testing.assert_close(net.fc.weight.detach(), ref.weight.detach(), rtol=1e-3, atol=2e-4)
    torch.testing.assert_close(net.fc.bias.detach(), ref.bias.detach(), rtol=1e-3, atol=2e-4)
    assert fopt.step_count == steps.shape[0]
    loss_sum, correct, samples = eng.metrics()
    assert samples == 1000 and correct == int(steps[:, 1].sum())
    # evaluation over the same data, in file order
    eng.evaluate(torch.arange(1000), 256)
    loss_sum, correct, samples = eng.metrics()
    with torch.no_grad():
        out = ref(_norm(images.cuda()))
        ref_loss = F.cross_entropy(out.double(), labels.long().cuda(), reduction='sum').item()
    assert samples == 1000
    assert abs(loss_sum - ref_loss) < 2e-3 * ref_loss
    assert abs(correct - int((out.argmax(1) == labels.long().cuda()).sum())) <= 2
    eng.close()


def test_engine_device_pool_steps():
    from pytorch_distributed_mnist_b200.ops.engine import NetEngine
    images, labels = _make(2048)
    net, fopt, ref, ropt = _build()
    eng = NetEngine(fopt.flat, fopt, max_batch=512)
    di, dl = images.cuda(), labels.cuda()
    ms = eng.train_device(di, dl, 512, 8)
    assert ms > 0 and fopt.step_count == 8
    for i in range(8):
        b = i % 4
        x, y = _norm(di[b * 512:(b + 1) * 512]), dl[b * 512:(b + 1) * 512].long()
        loss = F.cross_entropy(ref(x), y)
        ropt.zero_grad()
        loss.backward()
        ropt.step()
    torch.testing.assert_close(net.fc.weight.detach(), ref.weight.detach(), rtol=1e-3, atol=2e-4)
    eng.close()


def test_run_train_resume_evaluate_single_gpu(tmp_path, free_port, capsys):
    """BASELINE config 5 tail: --evaluate --resume on 1 GPU, through the public CLI path."""
    from pytorch_distributed_mnist_b200.cli import build_parser
    from pytorch_distributed_mnist_b200 import runner
    chk = str(tmp_path / 'checkpoints')
    base = ['--world-size', '1', '--backend', 'gloo', '--synthetic', '--train-size', '6000', '--test-size', '2000',
            '--batch-size', '64', '-i', 'tcp://127.0.0.1:{}'.format(free_port), '--checkpoint-dir', chk,
            '--device', 'cuda']
    args = build_parser().parse_args(base + ['--epochs', '2'])
    args.rank = 0
    runner.best_acc = 0
    out = runner.run(args)
    hist = out['history']
    assert len(hist) == 2 and hist[1]['train_loss'] < hist[0]['train_loss']
    assert hist[1]['test_acc'] > 0.85
    assert os.path.isfile(os.path.join(chk, 'checkpoint_1.pth.tar'))
    assert os.path.isfile(os.path.join(chk, 'model_best.pth.tar'))
    text = capsys.readouterr().out
    assert 'Epoch: 1/2,' in text and 'train loss:' in text and '%' in text

    # evaluate from the checkpoint: same numbers as the last epoch's test pass
    args = build_parser().parse_args(base + ['--evaluate', '--resume', os.path.join(chk, 'checkpoint_1.pth.tar')])
    args.rank = 0
    ev = runner.run(args)
    assert abs(ev['test_acc'] - hist[1]['test_acc']) < 1e-9
    assert abs(ev['test_loss'] - hist[1]['test_loss']) < 1e-6

    # the checkpoint loads into the stock stack (reference layout)
    ck = torch.load(os.path.join(chk, 'checkpoint_1.pth.tar'))
    assert set(ck.keys()) == {'epoch', 'state_dict', 'best_acc', 'optimizer'}
    assert list(ck['state_dict'].keys()) == ['module.fc.weight', 'module.fc.bias']
    lin = torch.nn.Linear(784, 10)
    lin.load_state_dict({k.replace('module.fc.', ''): v for k, v in ck['state_dict'].items()})
    stock = torch.optim.Adam(lin.parameters())
    stock.load_state_dict(ck['optimizer'])
    assert float(stock.state[lin.weight]['step']) == ck['optimizer']['state'][0]['step'].item() > 0


def test_fused_and_generic_paths_agree(free_port):
    """Same data order through the fused engine and through the autograd loop."""
    import torch.distributed as dist
    from pytorch_distributed_mnist_b200 import (Net, FusedAdam, Trainer, MNISTDataLoader, DistributedDataParallel)
    dist.init_process_group('gloo', init_method='tcp://127.0.0.1:{}'.format(free_port), world_size=1, rank=0)
    try:
        results = []
        for fused in ('on', 'off'):
            torch.manual_seed(0)
            model = DistributedDataParallel(Net().cuda(), device_ids=[0])
            opt = FusedAdam(model.parameters(), lr=1e-3)
            tl = MNISTDataLoader('nowhere', 64, 2, train=True, synthetic=True, size=2048)
            vl = MNISTDataLoader('nowhere', 64, 2, train=False, synthetic=True, size=512)
            tr = Trainer(model, opt, tl, vl, torch.device('cuda', 0), fused=fused)
            assert (tr.engine is not None) == (fused == 'on')
            tl.set_sample_epoch(0)
            a, b = tr.train()
            c, d = tr.evaluate()
            results.append((a.average, b.accuracy, c.average, d.accuracy,
                            model.module.fc.weight.detach().clone()))
            tr.close()
            model.close()
        f, g = results
        assert abs(f[0] - g[0]) < 2e-3 and abs(f[1] - g[1]) < 5e-3
        assert abs(f[2] - g[2]) < 2e-3 and abs(f[3] - g[3]) < 5e-3
        torch.testing.assert_close(f[4], g[4], rtol=1e-3, atol=2e-4)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('placement', ['device', 'zerocopy'])
@pytest.mark.parametrize('kernel', ['per_step', 'persistent'])
def test_in_kernel_gather_matches_staged_path(placement, kernel):
    """Same index plan through the pinned-ring path and through the in-kernel gather (dataset in
    HBM / read from pinned host memory over PCIe).  With an explicit CTA count both passes run the
    per-step FMA kernel and must agree bit for bit; by default the gather pass is ONE launch of the
    persistent tensor-core kernel (different summation order: fp32-level agreement)."""
    from pytorch_distributed_mnist_b200.ops import _lib
    from pytorch_distributed_mnist_b200.ops.engine import NetEngine
    images, labels = _make(3000)
    idx = torch.randperm(3000, generator=torch.Generator().manual_seed(5))[:2900]
    ctas = int(_lib.lib().pdm_net_default_ctas(128)) if kernel == 'per_step' else 0
    outs = []
    for mode in ('staged', placement):
        net, fopt, _, _ = _build()
        eng = NetEngine(fopt.flat, fopt, max_batch=128, slots=4, workers=2)
        eng.set_dataset(images, labels)
        eng.place_dataset(mode)
        before = eng.launches
        steps, ms = eng.train(idx, 128, ctas=ctas)
        assert ms > 0 and steps.shape == (23, 2)
        outs.append((steps.copy(), fopt.flat.params.clone(), eng.launches - before))
        eng.close()
    if kernel == 'per_step' or _lib.lib().pdm_net_tc_supported(0) != 1:
        assert np.array_equal(outs[0][0], outs[1][0])
        assert torch.equal(outs[0][1], outs[1][1])
    else:
        assert outs[0][2] == 23 and outs[1][2] == 1
        np.testing.assert_allclose(outs[0][0][:, 0], outs[1][0][:, 0], rtol=1e-4, atol=1e-4)
        assert np.abs(outs[0][0][:, 1] - outs[1][0][:, 1]).max() <= 1
        torch.testing.assert_close(outs[0][1], outs[1][1], rtol=1e-3, atol=1e-4)
