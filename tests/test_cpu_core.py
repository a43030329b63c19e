"""CPU checks of the framework shell against the reference's own classes."""
import math
import os

import numpy as np
import pytest
import torch

import pytorch_distributed_mnist_b200 as pdm
from pytorch_distributed_mnist_b200.cli import build_parser, detect_launcher, resolve_local_rank
from pytorch_distributed_mnist_b200.data import mnist
from pytorch_distributed_mnist_b200.data.loader import ShardSampler
from pytorch_distributed_mnist_b200.parallel.ddp import assign_buckets


# ---------------------------------------------------------------- CLI
REFERENCE_DEFAULTS = dict(root='data', workers=4, epochs=20, start_epoch=0, batch_size=256, lr=1e-3, momentum=0.9,
                          weight_decay=1e-4, resume='', evaluate=False, backend='nccl',
                          init_method='tcp://127.0.0.1:23456', world_size=1, rank=0, seed=None)


def test_cli_defaults_match_reference():
    args = build_parser().parse_args([])
    for k, v in REFERENCE_DEFAULTS.items():
        assert getattr(args, k) == v, k


def test_cli_short_and_long_spellings():
    a = build_parser().parse_args(['-j', '8', '-e', '-i', 'tcp://127.0.0.1:1', '-s', '4', '-r', '2',
                                   '--learning-rate', '0.5', '--weight-decay', '0.25', '--start-epoch', '3'])
    assert (a.workers, a.evaluate, a.init_method, a.world_size, a.rank) == (8, True, 'tcp://127.0.0.1:1', 4, 2)
    assert (a.lr, a.weight_decay, a.start_epoch) == (0.5, 0.25, 3)


@pytest.mark.parametrize('flag', ['--local-rank=3', '--local_rank=3'])
def test_local_rank_dash_and_underscore(flag, monkeypatch):
    """torch >= 2.0 launch passes --local-rank; the reference only declares --local_rank."""
    monkeypatch.delenv('LOCAL_RANK', raising=False)
    monkeypatch.delenv('RANK', raising=False)
    a = build_parser().parse_args([flag])
    assert a.local_rank == 3
    assert detect_launcher(a, [flag]) == 'launch'


def test_launcher_detection(monkeypatch):
    monkeypatch.delenv('LOCAL_RANK', raising=False)
    monkeypatch.delenv('RANK', raising=False)
    monkeypatch.delenv('TORCHELASTIC_RUN_ID', raising=False)
    a = build_parser().parse_args([])
    assert detect_launcher(a, []) == 'spawn'
    monkeypatch.setenv('LOCAL_RANK', '5')
    assert detect_launcher(a, []) == 'launch'
    assert resolve_local_rank(a) == 5
    a = build_parser().parse_args(['--launcher', 'spawn'])
    assert detect_launcher(a, []) == 'spawn'


# ---------------------------------------------------------------- metrics / schedule
def test_metrics_match_reference(reference_module):
    ours, ref = pdm.Average(), reference_module.Average()
    for v, n in [(0.5, 64), (1.25, 24), (3.0, 1)]:
        ours.update(v, n)
        ref.update(v, n)
    assert str(ours) == str(ref) and ours.average == ref.average
    torch.manual_seed(0)
    oa, ra = pdm.Accuracy(), reference_module.Accuracy()
    for _ in range(3):
        out, tgt = torch.randn(17, 10), torch.randint(0, 10, (17,))
        oa.update(out, tgt)
        ra.update(out, tgt)
    assert str(oa) == str(ra) and oa.accuracy == ra.accuracy
    with pytest.raises(ZeroDivisionError):
        pdm.Average().average


def test_lr_schedule_matches_reference(reference_module):
    class A:
        lr = 1e-3
    lin = torch.nn.Linear(2, 2)
    o1 = torch.optim.Adam(lin.parameters(), lr=A.lr)
    o2 = torch.optim.Adam(lin.parameters(), lr=A.lr)
    for epoch in [0, 1, 9, 10, 19, 20, 35]:
        pdm.adjust_learning_rate(o1, epoch, A)
        reference_module.adjust_learning_rate(o2, epoch, A)
        assert o1.param_groups[0]['lr'] == o2.param_groups[0]['lr']


# ---------------------------------------------------------------- model
def test_net_matches_reference(reference_module):
    torch.manual_seed(0)
    ours = pdm.Net()
    ref = reference_module.Net()
    ref.load_state_dict(ours.state_dict())
    assert list(ours.state_dict().keys()) == list(ref.state_dict().keys()) == ['fc.weight', 'fc.bias']
    assert sum(p.numel() for p in ours.parameters()) == 7850
    x = torch.randn(5, 1, 28, 28)
    assert torch.equal(ours(x), ref(x))


def test_cnn_shape_and_param_count():
    m = pdm.CNN()
    assert sum(p.numel() for p in m.parameters()) == 1199882
    assert m(torch.randn(3, 1, 28, 28)).shape == (3, 10)
    assert pdm.build_model('cnn', 2).fc1.in_features == 128 * 144


# ---------------------------------------------------------------- data
def test_idx_roundtrip_and_torchvision_reads_it(tmp_path):
    root = str(tmp_path)
    mnist.write_idx_tree(root, seed=7, train_size=300, test_size=100)
    images, labels, src = mnist.load_mnist(root, train=True)
    assert src.startswith('idx:') and images.shape == (300, 784) and labels.shape == (300,)
    exp_i, exp_l = mnist.synthetic_mnist(True, seed=7, size=300)
    assert np.array_equal(images.reshape(300, 28, 28), exp_i) and np.array_equal(labels, exp_l)
    from torchvision import datasets
    ds = datasets.MNIST(root, train=False, download=True)          # must not try the network
    assert len(ds) == 100


def test_loader_matches_reference_loader(tmp_path, reference_module):
    """Same bytes through the reference's MNISTDataLoader (PIL + ToTensor + Normalize per sample)
    and ours (batched uint8 gather): identical tensors, bit for bit."""
    root = str(tmp_path)
    mnist.write_idx_tree(root, seed=3, train_size=200, test_size=130)
    ref = reference_module.MNISTDataLoader(root, 64, num_workers=0, train=False)
    ours = pdm.MNISTDataLoader(root, 64, num_workers=0, train=False)
    assert len(ref) == len(ours) == 3
    for (rx, ry), (ox, oy) in zip(ref, ours):
        assert rx.shape == ox.shape and torch.equal(rx, ox)
        assert ry.dtype == oy.dtype == torch.int64 and torch.equal(ry, oy)


@pytest.mark.parametrize('n,world', [(60000, 4), (1001, 3), (10, 8), (7, 2)])
def test_sampler_matches_distributed_sampler(n, world):
    from torch.utils.data import DistributedSampler
    ds = list(range(n))
    for rank in range(world):
        for epoch in (0, 1, 5):
            ref = DistributedSampler(ds, num_replicas=world, rank=rank)
            ref.set_epoch(epoch)
            ours = ShardSampler(n, num_replicas=world, rank=rank)
            ours.set_epoch(epoch)
            assert list(ref) == ours.indices().tolist()
            assert len(ref) == len(ours)


def test_steps_per_epoch_defaults_w4():
    """defaults at W=4 give B=64, 15,000 samples, 235 steps, last batch 24."""
    s = ShardSampler(60000, num_replicas=4, rank=1)
    assert len(s) == 15000
    assert math.ceil(15000 / 64) == 235 and 15000 - 234 * 64 == 24


def test_synthetic_set_is_fixed_and_learnable():
    a = mnist.synthetic_mnist(True, seed=1234, size=500)
    b = mnist.synthetic_mnist(True, seed=1234, size=500)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    assert a[0].dtype == np.uint8 and a[0].shape == (500, 28, 28) and set(np.unique(a[1])) <= set(range(10))


# ---------------------------------------------------------------- buckets (SURVEY 2.5 bucketing facts)
def test_bucket_assignment():
    assert assign_buckets([7840, 10]) == [[1, 0]]                      # Net: one 31,400-byte bucket
    cnn = [p.numel() for p in pdm.CNN().parameters()]
    buckets = assign_buckets(cnn)
    assert [i for b in buckets for i in b] == list(reversed(range(len(cnn))))
    assert all(b == sorted(b, reverse=True) for b in buckets)
    first_bytes = 4 * sum(cnn[i] for i in buckets[0])
    assert first_bytes >= 1024 * 1024 and len(buckets) >= 2
    big = assign_buckets([30 * 1024 * 1024 // 4] * 3, cap_bytes=25 * 1024 * 1024)
    assert len(big) == 3


# ---------------------------------------------------------------- optimizers on CPU
@pytest.mark.parametrize('kind', ['adam', 'sgd'])
def test_fused_optimizers_cpu_match_torch(kind):
    torch.manual_seed(0)
    net, ref = pdm.Net(), pdm.Net()
    ref.load_state_dict(net.state_dict())
    if kind == 'adam':
        ours, stock = pdm.FusedAdam(net.parameters(), lr=1e-2), torch.optim.Adam(ref.parameters(), lr=1e-2)
    else:
        ours = pdm.FusedSGD(net.parameters(), lr=1e-2, momentum=0.9, weight_decay=1e-4)
        stock = torch.optim.SGD(ref.parameters(), lr=1e-2, momentum=0.9, weight_decay=1e-4)
    for _ in range(5):
        x, y = torch.randn(16, 784), torch.randint(0, 10, (16,))
        for m, o in ((net, ours), (ref, stock)):
            o.zero_grad()
            torch.nn.functional.cross_entropy(m(x), y).backward()
            o.step()
    torch.testing.assert_close(net.fc.weight, ref.fc.weight, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(net.fc.bias, ref.fc.bias, rtol=1e-5, atol=1e-6)


def test_adam_state_dict_is_stock_layout_and_accepts_old_format():
    torch.manual_seed(0)
    net = pdm.Net()
    opt = pdm.FusedAdam(net.parameters(), lr=1e-3)
    torch.nn.functional.cross_entropy(net(torch.randn(4, 784)), torch.tensor([1, 2, 3, 4])).backward()
    opt.step()
    sd = opt.state_dict()
    stock = torch.optim.Adam(pdm.Net().parameters())
    stock.load_state_dict(sd)                                           # ours -> stock
    assert set(sd['param_groups'][0].keys()) == set(stock.state_dict()['param_groups'][0].keys())
    assert sd['state'][0]['step'].dtype == torch.float32 and sd['state'][0]['step'].item() == 1.0
    assert sd['state'][0]['exp_avg'].shape == (10, 784) and sd['state'][1]['exp_avg_sq'].shape == (10,)
    # old torch wrote `step` as a Python int and fewer param_group keys
    old = {'state': {0: {'step': 7, 'exp_avg': torch.ones(10, 784), 'exp_avg_sq': torch.ones(10, 784)},
                     1: {'step': 7, 'exp_avg': torch.ones(10), 'exp_avg_sq': torch.ones(10)}},
           'param_groups': [{'lr': 0.01, 'betas': (0.9, 0.999), 'eps': 1e-8, 'weight_decay': 0, 'amsgrad': False,
                             'params': [0, 1]}]}
    opt.load_state_dict(old)
    assert opt.step_count == 7 and opt.param_groups[0]['lr'] == 0.01 and 'capturable' in opt.param_groups[0]
    assert float(opt.flat.exp_avg[:7850].min()) == 1.0


# ---------------------------------------------------------------- native library builds here (no GPU needed)
def test_native_library_builds_and_exports():
    from pytorch_distributed_mnist_b200.ops import _lib
    lib = _lib.lib()
    assert lib.pdm_version() >= 100
    for sym in ('pdm_net_train_step', 'pdm_net_eval_step', 'pdm_comm_allreduce_optim', 'pdm_engine_run',
                'pdm_optim_step', 'pdm_cross_entropy'):
        assert hasattr(lib, sym)
    if not torch.cuda.is_available():
        assert lib.pdm_device_ok(0) == 0


def test_resume_auto_picks_newest_checkpoint(tmp_path):
    from pytorch_distributed_mnist_b200 import checkpoint as ckpt
    assert ckpt.latest_checkpoint(str(tmp_path / 'missing')) is None
    assert ckpt.latest_checkpoint(str(tmp_path)) is None
    for name in ('checkpoint_2.pth.tar', 'checkpoint_10.pth.tar', 'checkpoint_9.pth.tar', 'model_best.pth.tar',
                 'checkpoint_11.pth.tar.tmp'):
        (tmp_path / name).write_bytes(b'x')
    assert ckpt.latest_checkpoint(str(tmp_path)) == str(tmp_path / 'checkpoint_10.pth.tar')


def test_ddp_reports_parameters_without_gradient():
    """A parameter that never gets a gradient leaves its bucket unreduced; like stock DDP this must
    be an error at the next forward, not silent divergence."""
    import pytorch_distributed_mnist_b200 as pdm

    class TwoHeads(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.a = torch.nn.Linear(4, 3)
            self.b = torch.nn.Linear(4, 3)

        def forward(self, x, use_b=True):
            return self.a(x) + (self.b(x) if use_b else 0)

    model = pdm.DistributedDataParallel(TwoHeads())
    opt = pdm.FusedSGD(model.parameters(), lr=0.1)
    x = torch.randn(5, 4)
    model(x).sum().backward()          # complete iteration
    opt.step()
    model(x).sum().backward()          # two forwards in a row are fine when each backward completes
    opt.step()
    model(x, use_b=False).sum().backward()      # head b gets no gradient
    with pytest.raises(RuntimeError, match='did not produce gradients for every parameter'):
        model(x)


def test_epoch_plan_is_built_at_set_sample_epoch():
    """set_sample_epoch (ref:231) prepares the epoch's index plan; train() only consumes it.  The plan handed
    out must be exactly what the sampler yields for that epoch, and a stale plan is never used."""
    from pytorch_distributed_mnist_b200.data.loader import MNISTDataLoader, ShardSampler
    loader = MNISTDataLoader('unused', 32, train=True, synthetic=True, size=1000, samples_per_epoch=1500)
    loader.sampler = ShardSampler(1000, num_replicas=2, rank=1)
    loader.set_sample_epoch(3)
    assert loader._plan is not None and loader._plan[0] == 3
    got = loader.epoch_indices()
    want = torch.cat([loader.sampler.indices(3), loader.sampler.indices(4), loader.sampler.indices(5)])[:1500]
    assert torch.equal(got, want)
    assert torch.equal(loader.epoch_indices(), want)          # asking twice in one epoch is fine
    loader.sampler.set_epoch(7)                                # epoch moved behind the loader's back
    assert torch.equal(loader.epoch_indices()[:500], loader.sampler.indices(7))


def test_native_reducer_launches_buckets_in_index_order():
    """bucket-ready counting and in-order launches live in csrc/reducer.cu.  A later bucket
    that becomes ready first is held back until every earlier one has gone out."""
    from pytorch_distributed_mnist_b200.parallel.reducer import NativeReducer
    # 5 parameters; bucket 0 = {4, 3}, bucket 1 = {2}, bucket 2 = {1, 0} (reverse order, like assign_buckets)
    r = NativeReducer([2, 2, 1, 0, 0], [(32, 48), (16, 32), (0, 16)])
    assert r.arm()
    assert r.mark_ready(2) == (0, 0)            # bucket 1 complete, bucket 0 still open: held
    assert r.mark_ready(4) == (0, 0)
    assert r.mark_ready(3) == (0, 2)            # bucket 0 closes: 0 and the held 1 go out together
    assert r.mark_ready(0) == (2, 0)
    st = r.state()
    assert st['next'] == 2 and st['pending'] == [0, 0, 1] and st['order'] == [0, 1]
    assert not r.arm()                          # unfinished reduction is refused, state kept
    assert r.state()['pending'] == [0, 0, 1]
    assert r.mark_ready(1) == (2, 1)
    assert r.state()['launches'] == 3
    with pytest.raises(RuntimeError, match='marked twice'):
        r.mark_ready(1)
    assert r.arm() and r.state()['order'] == []
    r.close()
    with pytest.raises(RuntimeError, match='4-float aligned'):
        NativeReducer([0], [(0, 6)])
    with pytest.raises(RuntimeError, match='maps to bucket'):
        NativeReducer([1], [(0, 8)])


def test_flat_state_relayout_moves_parameters_gradients_and_moments():
    """FlatState.relayout (used by the reducer's bucket rebuild): values follow their parameter into the new
    memory order, every view is re-created, training continues exactly as without the move."""
    from pytorch_distributed_mnist_b200.parallel.flat import FlatState

    def make():
        torch.manual_seed(3)
        return torch.nn.Sequential(torch.nn.Linear(7, 5), torch.nn.Linear(5, 3), torch.nn.Linear(3, 2))

    a, b = make(), make()
    oa = pdm.FusedAdam(a.parameters(), lr=1e-2)
    ob = pdm.FusedAdam(b.parameters(), lr=1e-2)
    x, y = torch.randn(6, 7), torch.randint(0, 2, (6,))

    def step(m, o):
        o.zero_grad()
        torch.nn.functional.cross_entropy(m(x), y).backward()
        o.step()

    step(a, oa); step(b, ob)
    flat = ob.flat
    assert isinstance(flat, FlatState) and flat.layout == list(range(6))
    flat.relayout([4, 5, 2, 3, 0, 1])                       # last layer first
    assert flat.offsets[4] == 0 and flat.layout == [4, 5, 2, 3, 0, 1]
    first = list(b.parameters())[4]
    assert first.data_ptr() == flat.params.data_ptr() and first.grad.data_ptr() == flat.grads.data_ptr()
    for p, q in zip(a.parameters(), b.parameters()):
        assert torch.equal(p, q)
    step(a, oa); step(b, ob)                                # Adam moments moved too, or this diverges
    for p, q in zip(a.parameters(), b.parameters()):
        torch.testing.assert_close(p, q, rtol=0, atol=0)
    sa, sb = oa.state_dict()['state'], ob.state_dict()['state']
    for k in sa:
        assert torch.equal(sa[k]['exp_avg'], sb[k]['exp_avg']) and torch.equal(sa[k]['exp_avg_sq'], sb[k]['exp_avg_sq'])
    with pytest.raises(ValueError):
        flat.relayout([0, 1, 2])
    flat.layout_frozen = True
    with pytest.raises(RuntimeError, match='native engine'):
        flat.relayout(list(range(6)))
