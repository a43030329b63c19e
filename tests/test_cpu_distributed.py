"""Multi-process CPU/gloo plumbing: DDP replacement vs stock DDP, both launch modes,
checkpoint round trips against the reference's own code (BASELINE config 1)."""
import os
import subprocess
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TRAIN = os.path.join(ROOT, 'train.py')


def _env():
    env = dict(os.environ)
    env['PYTHONPATH'] = ROOT + os.pathsep + env.get('PYTHONPATH', '')
    for k in ('LOCAL_RANK', 'RANK', 'WORLD_SIZE', 'MASTER_ADDR', 'MASTER_PORT', 'TORCHELASTIC_RUN_ID'):
        env.pop(k, None)
    return env


def _ddp_worker(rank, world, port, out_dir):
    import pytorch_distributed_mnist_b200 as pdm
    dist.init_process_group('gloo', init_method='tcp://127.0.0.1:{}'.format(port), world_size=world, rank=rank)
    try:
        torch.manual_seed(100 + rank)                      # different init per rank: broadcast must fix it
        ours = pdm.DistributedDataParallel(pdm.CNN())
        torch.manual_seed(100)                             # rank 0's init
        stock_mod = pdm.CNN()
        stock = torch.nn.parallel.DistributedDataParallel(stock_mod)
        for a, b in zip(ours.module.parameters(), stock_mod.parameters()):
            assert torch.equal(a, b), 'initial broadcast from rank 0 failed'
        assert len(ours.buckets) >= 2
        o_opt = pdm.FusedAdam(ours.parameters(), lr=1e-3)
        s_opt = torch.optim.Adam(stock.parameters(), lr=1e-3)
        ours.eval(); stock.eval()                          # dropout off: deterministic comparison
        g = torch.Generator().manual_seed(7 + rank)
        for _ in range(3):
            x = torch.randn(8, 1, 28, 28, generator=g)
            y = torch.randint(0, 10, (8,), generator=g)
            for m, o in ((ours, o_opt), (stock, s_opt)):
                loss = torch.nn.functional.cross_entropy(m(x), y)
                o.zero_grad()
                loss.backward()
                o.step()
        for a, b in zip(ours.module.parameters(), stock_mod.parameters()):
            torch.testing.assert_close(a, b, rtol=1e-4, atol=1e-6)
        assert list(ours.state_dict().keys()) == list(stock.state_dict().keys())
        assert ours.stats['bucket_launches'] == 3 * len(ours.buckets)
        torch.save({k: v for k, v in ours.state_dict().items()}, os.path.join(out_dir, 'rank{}.pt'.format(rank)))
        ours.close()
    finally:
        dist.destroy_process_group()


def test_ddp_replacement_matches_stock_ddp_gloo(tmp_path, free_port):
    mp.spawn(_ddp_worker, args=(2, free_port, str(tmp_path)), nprocs=2)
    a = torch.load(str(tmp_path / 'rank0.pt'))
    b = torch.load(str(tmp_path / 'rank1.pt'))
    for k in a:
        assert torch.equal(a[k], b[k]), 'replicas diverged: ' + k


def _mismatch_worker(rank, world, port):
    import pytorch_distributed_mnist_b200 as pdm
    dist.init_process_group('gloo', init_method='tcp://127.0.0.1:{}'.format(port), world_size=world, rank=rank)
    try:
        model = pdm.Net() if rank == 0 else pdm.CNN()
        with pytest.raises(RuntimeError, match='parameter shapes differ'):
            pdm.DistributedDataParallel(model)
    finally:
        dist.destroy_process_group()


def test_ddp_detects_mismatched_models(free_port):
    mp.spawn(_mismatch_worker, args=(2, free_port), nprocs=2)


COMMON = ['--backend', 'gloo', '--synthetic', '--train-size', '2048', '--test-size', '512', '--batch-size', '128',
          '--workers', '3']


def test_spawn_mode_train_resume_evaluate(tmp_path, free_port):
    """way 2 (README:33) + resume (README:50-52) + evaluate (README:55-60) at W=2 on CPU."""
    cwd = str(tmp_path)
    url = 'tcp://127.0.0.1:{}'.format(free_port)
    r = subprocess.run([sys.executable, TRAIN, '--world-size', '2', '--epochs', '2', '-i', url] + COMMON,
                       cwd=cwd, env=_env(), capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stdout + r.stderr
    out = r.stdout
    assert out.count('rank: ') == 2 and 'rank: 1, device count: ' in out and 'workers:2' in out   # ceil(3/2)
    assert out.count('Epoch: 1/2,') == 2                                  # every rank prints
    assert os.path.isfile(os.path.join(cwd, 'checkpoints', 'checkpoint_0.pth.tar'))
    assert os.path.isfile(os.path.join(cwd, 'checkpoints', 'checkpoint_1.pth.tar'))
    assert os.path.isfile(os.path.join(cwd, 'checkpoints', 'model_best.pth.tar'))
    ck = torch.load(os.path.join(cwd, 'checkpoints', 'checkpoint_1.pth.tar'))
    assert ck['epoch'] == 2 and 0.0 <= ck['best_acc'] <= 1.0

    r = subprocess.run([sys.executable, TRAIN, '--world-size', '2', '--epochs', '3', '-i', url, '--resume',
                        'checkpoints/checkpoint_1.pth.tar'] + COMMON,
                       cwd=cwd, env=_env(), capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "=> loaded checkpoint 'checkpoints/checkpoint_1.pth.tar' (epoch 2)" in r.stdout
    assert 'Epoch: 2/3,' in r.stdout and 'Epoch: 1/3,' not in r.stdout

    r = subprocess.run([sys.executable, TRAIN, '--world-size', '1', '-i', url, '--evaluate', '--resume',
                        'checkpoints/model_best.pth.tar'] + COMMON,
                       cwd=cwd, env=_env(), capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stdout + r.stderr
    assert 'test loss: ' in r.stdout and 'test acc: ' in r.stdout and '%.' in r.stdout

    r = subprocess.run([sys.executable, TRAIN, '--world-size', '1', '-i', url, '--epochs', '0', '--resume',
                        'nope.pth.tar'] + COMMON, cwd=cwd, env=_env(), capture_output=True, text=True, timeout=240)
    assert r.returncode == 0 and "=> no checkpoint found at 'nope.pth.tar'" in r.stdout


def test_launch_mode_torch_distributed_launch(tmp_path, free_port):
    """way 1 (README:19): torch.distributed.launch passes --local-rank=N."""
    import socket
    s = socket.socket(); s.bind(('127.0.0.1', 0)); mport = s.getsockname()[1]; s.close()
    cmd = [sys.executable, '-m', 'torch.distributed.launch', '--nproc_per_node=2', '--master_addr', '127.0.0.1',
           '--master_port', str(mport), TRAIN, '--world-size', '2', '--epochs', '1',
           '-i', 'tcp://127.0.0.1:{}'.format(free_port)] + COMMON
    r = subprocess.run(cmd, cwd=str(tmp_path), env=_env(), capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.count('Epoch: 0/1,') == 2
    assert os.path.isfile(os.path.join(str(tmp_path), 'checkpoints', 'checkpoint_0.pth.tar'))


def _ref_ckpt_worker(rank, world, port, path, ref_path):
    """The SURVEY probe recipe: reference save_checkpoint + stock DDP + stock Adam."""
    import importlib.util
    spec = importlib.util.spec_from_file_location('ref_mod', ref_path)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    dist.init_process_group('gloo', init_method='tcp://127.0.0.1:{}'.format(port), world_size=world, rank=rank)
    try:
        torch.manual_seed(0)
        model = torch.nn.parallel.DistributedDataParallel(ref.Net())
        opt = torch.optim.Adam(model.parameters(), lr=1e-3)
        for _ in range(3):
            loss = torch.nn.functional.cross_entropy(model(torch.randn(16, 1, 28, 28)), torch.randint(0, 10, (16,)))
            opt.zero_grad(); loss.backward(); opt.step()
        os.chdir(path)
        ref.save_checkpoint({'epoch': 4, 'state_dict': model.state_dict(), 'best_acc': 0.5,
                             'optimizer': opt.state_dict()}, True, 3)
    finally:
        dist.destroy_process_group()


def test_reference_checkpoint_resumes_here_and_ours_loads_in_stock(tmp_path, free_port):
    ref_path = '/root/reference/multi_proc_single_gpu.py'
    if not os.path.isfile(ref_path):
        ref_path = os.path.join(ROOT, 'baseline', '_ref', 'multi_proc_single_gpu.py')
    if not os.path.isfile(ref_path):
        pytest.skip('reference module not available')
    mp.spawn(_ref_ckpt_worker, args=(1, free_port, str(tmp_path), ref_path), nprocs=1)
    ck_path = os.path.join(str(tmp_path), 'checkpoints', 'checkpoint_3.pth.tar')
    assert os.path.isfile(ck_path) and os.path.isfile(os.path.join(str(tmp_path), 'checkpoints', 'model_best.pth.tar'))
    # reference checkpoint -> our CLI: evaluate, then resume training
    r = subprocess.run([sys.executable, TRAIN, '--world-size', '1', '-i', 'tcp://127.0.0.1:{}'.format(free_port),
                        '--evaluate', '--resume', ck_path] + COMMON,
                       cwd=str(tmp_path), env=_env(), capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stdout + r.stderr
    assert 'best_acc: 0.5' in r.stdout and '(epoch 4)' in r.stdout and 'test acc: ' in r.stdout
    r = subprocess.run([sys.executable, TRAIN, '--world-size', '1', '-i', 'tcp://127.0.0.1:{}'.format(free_port),
                        '--epochs', '5', '--resume', ck_path] + COMMON,
                       cwd=str(tmp_path), env=_env(), capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stdout + r.stderr
    assert 'Epoch: 4/5,' in r.stdout
    # our checkpoint -> the stock stack (DDP(Net) + torch.optim.Adam), exactly as the reference would load it
    ours = torch.load(os.path.join(str(tmp_path), 'checkpoints', 'checkpoint_4.pth.tar'))
    theirs = torch.load(ck_path)
    assert ours['epoch'] == 5
    assert list(ours['state_dict'].keys()) == list(theirs['state_dict'].keys())
    assert ours['optimizer']['state'][0]['step'].item() == theirs['optimizer']['state'][0]['step'].item() + 16
    assert set(ours['optimizer']['param_groups'][0].keys()) == set(theirs['optimizer']['param_groups'][0].keys())
    lin = torch.nn.Linear(784, 10)
    lin.load_state_dict({k.replace('module.fc.', ''): v for k, v in ours['state_dict'].items()})
    torch.optim.Adam(lin.parameters()).load_state_dict(ours['optimizer'])


def _consolidate_worker(rank, world, port, out_dir):
    import torch.distributed as dist
    from pytorch_distributed_mnist_b200.parallel.flat import FlatState
    dist.init_process_group('gloo', init_method='tcp://127.0.0.1:{}'.format(port), world_size=world, rank=rank)
    try:
        torch.manual_seed(0)
        # one bucket large enough for the two-shot kernel (>= 256 KB), one small, odd sizes on purpose
        params = [torch.nn.Parameter(torch.zeros(70001)), torch.nn.Parameter(torch.zeros(33)), torch.nn.Parameter(torch.zeros(7))]
        flat = FlatState(params)
        lo, hi = flat.range_of(0, 0)
        small = flat.range_of(1, 2)
        truth_m = torch.arange(flat.numel, dtype=torch.float32) * 0.5
        truth_v = torch.arange(flat.numel, dtype=torch.float32) + 3.0
        # what the two-shot kernel leaves behind: only the owned slice of the big bucket is current
        flat.exp_avg.fill_(-1.0)
        flat.exp_avg_sq.fill_(-1.0)
        a, b = FlatState.slice_of(lo, hi, world, rank)
        flat.exp_avg[a:b] = truth_m[a:b]
        flat.exp_avg_sq[a:b] = truth_v[a:b]
        flat.exp_avg[small[0]:small[1]] = truth_m[small[0]:small[1]]       # small buckets: every rank updates all of it
        flat.exp_avg_sq[small[0]:small[1]] = truth_v[small[0]:small[1]]
        flat.mark_sharded(lo, hi, world)
        flat.mark_sharded(small[0], small[1], world)                        # below the threshold: not recorded
        assert list(flat.sharded) == [(lo, hi)] and flat.state_stale
        moved = flat.consolidate()
        assert moved == 2 * ((hi - lo) - (b - a)) and not flat.state_stale      # both moment buffers
        assert torch.equal(flat.exp_avg, truth_m) and torch.equal(flat.exp_avg_sq, truth_v)
        assert flat.consolidate() == 0                                      # nothing new since
        # the slices tile the bucket exactly, for every world size up to 8
        for w in range(1, 9):
            cuts = [FlatState.slice_of(lo, hi, w, r) for r in range(w)]
            assert cuts[0][0] == lo and cuts[-1][1] == hi and all(cuts[i][1] == cuts[i + 1][0] for i in range(w - 1))
        torch.save(1, os.path.join(out_dir, 'ok{}'.format(rank)))
    finally:
        dist.destroy_process_group()


def test_sharded_optimizer_state_is_consolidated(tmp_path, free_port):
    """Two-shot fused updates keep exp_avg/exp_avg_sq slices on their owners; consolidate() must
    rebuild the full state on every rank before rank 0 writes a checkpoint (ref:249-254)."""
    world = 3
    mp.spawn(_consolidate_worker, args=(world, free_port, str(tmp_path)), nprocs=world)
    assert all(os.path.isfile(str(tmp_path / 'ok{}'.format(r))) for r in range(world))


def _buffers_worker(rank, world, port, out_dir):
    import pytorch_distributed_mnist_b200 as pdm
    dist.init_process_group('gloo', init_method='tcp://127.0.0.1:{}'.format(port), world_size=world, rank=rank)
    try:
        torch.manual_seed(5)
        net = torch.nn.Sequential(torch.nn.Linear(6, 4), torch.nn.BatchNorm1d(4), torch.nn.Linear(4, 2))
        ours = pdm.DistributedDataParallel(net)
        opt = pdm.FusedSGD(ours.parameters(), lr=0.05)
        g = torch.Generator().manual_seed(30 + rank)        # different data per rank -> different running stats
        for _ in range(3):
            opt.zero_grad()
            ours(torch.randn(8, 6, generator=g)).sum().backward()
            opt.step()
        bn = net[1]
        rank0_mean = bn.running_mean.clone()
        dist.broadcast(rank0_mean, 0)                        # what rank 0 holds before the next forward
        if rank != 0:
            bn.running_mean.fill_(1000.0)                    # garbage that must be overwritten, not averaged in
        x = torch.randn(8, 6, generator=g)
        with torch.no_grad():
            batch_mean = net[0](x).mean(0)
        ours(x)
        want = (1 - bn.momentum) * rank0_mean + bn.momentum * batch_mean
        torch.testing.assert_close(bn.running_mean, want, rtol=1e-5, atol=1e-6)
        assert int(bn.num_batches_tracked) == 4
        torch.save(1, os.path.join(out_dir, 'r{}.pt'.format(rank)))
    finally:
        dist.destroy_process_group()


def test_buffers_are_rebroadcast_each_training_forward(tmp_path, free_port):
    """broadcast_buffers=True (stock default): BatchNorm statistics restart from rank 0's at every forward."""
    import pytorch_distributed_mnist_b200 as pdm
    mp.spawn(_buffers_worker, args=(2, free_port, str(tmp_path)), nprocs=2)
    assert os.path.isfile(str(tmp_path / 'r1.pt'))
    # single-process check of the mechanism itself: buffers are copied back through the flat view
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Linear(3, 3), torch.nn.BatchNorm1d(3))
    ddp = pdm.DistributedDataParallel(net)
    ddp(torch.randn(4, 3)).sum().backward()                 # world == 1: no-op path must not touch anything
    assert int(net[1].num_batches_tracked) == 1


class _TwoBranch(torch.nn.Module):
    """Each forward uses branch ``a`` or branch ``b``; ``never`` is not used at all."""

    def __init__(self):
        super().__init__()
        self.a = torch.nn.Linear(600, 500)        # 1.2 MB each: two buckets
        self.b = torch.nn.Linear(600, 500)
        self.c = torch.nn.Linear(500, 4)
        self.never = torch.nn.Linear(3, 3)

    def forward(self, x, use_b):
        return self.c(torch.relu(self.b(x) if use_b else self.a(x)))


def _unused_worker(rank, world, port, out_dir):
    import pytorch_distributed_mnist_b200 as pdm
    dist.init_process_group('gloo', init_method='tcp://127.0.0.1:{}'.format(port), world_size=world, rank=rank)
    try:
        torch.manual_seed(5)
        ours = pdm.DistributedDataParallel(_TwoBranch(), find_unused_parameters=True, bucket_cap_mb=1)
        torch.manual_seed(5)
        stock_mod = _TwoBranch()
        stock = torch.nn.parallel.DistributedDataParallel(stock_mod, find_unused_parameters=True)
        assert len(ours.buckets) >= 2
        o_opt = pdm.FusedSGD(ours.parameters(), lr=0.05)
        s_opt = torch.optim.SGD(stock.parameters(), lr=0.05)
        g = torch.Generator().manual_seed(11 + rank)
        for i in range(4):
            x = torch.randn(16, 600, generator=g)
            y = torch.randint(0, 4, (16,), generator=g)
            use_b = (i + rank) % 2 == 1                    # the ranks use different branches in the same step
            for m, o in ((ours, o_opt), (stock, s_opt)):
                o.zero_grad()
                torch.nn.functional.cross_entropy(m(x, use_b), y).backward()
                o.step()
        for (n, a), b in zip(ours.module.named_parameters(), stock_mod.parameters()):
            torch.testing.assert_close(a, b, rtol=1e-5, atol=1e-6, msg=lambda m, n=n: n + ': ' + m)
        st = ours._reducer.state()
        assert st['next'] == len(ours.buckets)             # the last pass reduced every bucket
        assert ours.stats['bucket_launches'] >= 4 * 2      # (the bucket rebuild after pass 1 may regroup them)
        assert ours.stats['unused_parameters'] == 4        # one branch (weight + bias) and `never`
        torch.save({k: v for k, v in ours.state_dict().items()}, os.path.join(out_dir, 'rank{}.pt'.format(rank)))
        ours.close()
    finally:
        dist.destroy_process_group()


def test_find_unused_parameters_matches_stock_ddp(tmp_path, free_port):
    """parameters without a gradient are counted as ready at the end of the backward pass by the
    native reducer; every bucket is still reduced on every rank although the ranks used different branches."""
    mp.spawn(_unused_worker, args=(2, free_port, str(tmp_path)), nprocs=2)
    a = torch.load(str(tmp_path / 'rank0.pt'))
    b = torch.load(str(tmp_path / 'rank1.pt'))
    for k in a:
        assert torch.equal(a[k], b[k]), 'replicas diverged: ' + k


class _OutOfOrder(torch.nn.Module):
    """Registered back to front: the layer used LAST comes first in ``parameters()``, so its bucket is
    complete long before the buckets in front of it and has to wait -- the case the rebuild fixes."""

    def __init__(self):
        super().__init__()
        self.last = torch.nn.Linear(520, 4)
        self.first = torch.nn.Linear(600, 520)       # 1.25 MB
        self.mid = torch.nn.Linear(520, 520)         # 1.08 MB

    def forward(self, x):
        return self.last(torch.relu(self.mid(torch.relu(self.first(x)))))


def _rebuild_worker(rank, world, port, out_dir):
    import pytorch_distributed_mnist_b200 as pdm
    dist.init_process_group('gloo', init_method='tcp://127.0.0.1:{}'.format(port), world_size=world, rank=rank)
    try:
        torch.manual_seed(9)
        ours = pdm.DistributedDataParallel(_OutOfOrder(), bucket_cap_mb=1)
        torch.manual_seed(9)
        stock_mod = _OutOfOrder()
        stock = torch.nn.parallel.DistributedDataParallel(stock_mod, bucket_cap_mb=1)
        names = [n for n, _ in ours.module.named_parameters()]
        before = [[names[i] for i in bk['params']] for bk in ours.buckets]
        assert before[-1] == ['last.bias', 'last.weight'] and len(before) == 3
        o_opt = pdm.FusedAdam(ours.parameters(), lr=1e-3)
        s_opt = torch.optim.Adam(stock.parameters(), lr=1e-3)
        g = torch.Generator().manual_seed(21 + rank)
        for i in range(4):
            x = torch.randn(16, 600, generator=g)
            y = torch.randint(0, 4, (16,), generator=g)
            for m, o in ((ours, o_opt), (stock, s_opt)):
                o.zero_grad()
                torch.nn.functional.cross_entropy(m(x), y).backward()
                o.step()
            if i == 0:
                assert ours._reducer.ready_order()[1], 'the bucket of the last layer should have been held back'
        assert ours.stats.get('rebuilt') == 1
        after = [[names[i] for i in bk['params']] for bk in ours.buckets]
        assert set(after[0]) >= {'last.bias', 'last.weight'} and after != before, after
        assert not ours._reducer.ready_order()[1], 'after the rebuild every bucket completes in launch order'
        for (n, a), b in zip(ours.module.named_parameters(), stock_mod.parameters()):
            torch.testing.assert_close(a, b, rtol=1e-4, atol=1e-6, msg=lambda m, n=n: n + ': ' + m)
        # the optimizer state moved with the parameters
        so, ss = o_opt.state_dict()['state'], s_opt.state_dict()['state']
        for k in ss:
            torch.testing.assert_close(so[k]['exp_avg'], ss[k]['exp_avg'], rtol=1e-4, atol=1e-7)
            torch.testing.assert_close(so[k]['exp_avg_sq'], ss[k]['exp_avg_sq'], rtol=1e-4, atol=1e-9)
        torch.save({k: v for k, v in ours.state_dict().items()}, os.path.join(out_dir, 'rank{}.pt'.format(rank)))
        ours.close()
    finally:
        dist.destroy_process_group()


def test_bucket_rebuild_after_the_first_iteration(tmp_path, free_port):
    """SURVEY N4 (Reducer::rebuild_buckets): a model whose parameters become ready in another order than
    the reverse registration order is regrouped after the first pass -- parameters, gradients and Adam
    moments move inside the flat buffer -- and keeps matching stock DDP + Adam."""
    mp.spawn(_rebuild_worker, args=(2, free_port, str(tmp_path)), nprocs=2)
    a = torch.load(str(tmp_path / 'rank0.pt'))
    b = torch.load(str(tmp_path / 'rank1.pt'))
    for k in a:
        assert torch.equal(a[k], b[k]), 'replicas diverged: ' + k
