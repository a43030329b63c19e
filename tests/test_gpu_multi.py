"""Multi-GPU checks of the peer-memory communicator: run with
`pytest -m gpu` on a box with >= 2 GPUs (gpurun --gpus 2); skipped otherwise.

One process per GPU (mp.spawn), NCCL process group for the plumbing and as the oracle."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]

MEAN, STD = 0.1307, 0.3081


def _world():
    n = torch.cuda.device_count()
    return 8 if n >= 8 else (4 if n >= 4 else 2)


def _init(rank, world, port):
    torch.cuda.set_device(rank)
    dist.init_process_group('nccl', init_method='tcp://127.0.0.1:{}'.format(port), world_size=world, rank=rank,
                            device_id=torch.device('cuda', rank))


def _rand(seed, n):
    return torch.randn(n, generator=torch.Generator().manual_seed(seed))


def _comm_worker(rank, world, port):
    from pytorch_distributed_mnist_b200.ops import _lib
    from pytorch_distributed_mnist_b200.parallel.peer import PeerComm
    _init(rank, world, port)
    try:
        assert PeerComm.supported(world)
        dev = torch.device('cuda', rank)
        sizes = [4, 1024, 7852, 300000, 1 << 20]
        nmax = max(sizes)
        comm = PeerComm(rank, world, dev, data_bytes=nmax * 4 * 5 + 8192)
        a, b, p, m, v = (comm.data_tensor(nmax) for _ in range(5))
        comm.barrier()
        # broadcast from every root
        for root in range(world):
            a.copy_(_rand(100 + rank, nmax))
            torch.cuda.synchronize()
            comm.broadcast(a, root=root)
            torch.cuda.synchronize()
            assert torch.equal(a.cpu(), _rand(100 + root, nmax)), 'broadcast root {}'.format(root)
        # plain all-reduce, both algorithms, bitwise vs the rank-ordered fp32 sum
        for n in sizes:
            for algo in (1, 2):
                for rep in range(2):
                    a[:n].copy_(_rand(1000 * rep + 10 * n % 997 + rank, n))
                    b[:n].zero_()
                    torch.cuda.synchronize()
                    dst = b[:n] if algo == 1 else a[:n]
                    comm.allreduce(a[:n], dst, average=True, algo=algo)
                    torch.cuda.synchronize()
                    exp = torch.zeros(n)
                    for r in range(world):
                        exp = exp + _rand(1000 * rep + 10 * n % 997 + r, n)
                    exp = exp * (1.0 / world)
                    assert torch.equal(dst.cpu(), exp), 'allreduce n={} algo={} rep={}'.format(n, algo, rep)
        # fused all-reduce + Adam: vs torch on the averaged gradient; replicas bit-identical
        for n, algo in ((7852, 1), (300000, 2), (1 << 20, 2)):
            p[:n].copy_(_rand(5, n))
            m[:n].zero_()
            v[:n].zero_()
            ref_p = _rand(5, n).cuda().requires_grad_(True)
            ropt = torch.optim.Adam([ref_p], lr=1e-2)
            for t in range(1, 4):
                a[:n].copy_(_rand(7000 + 10 * t + rank, n))
                torch.cuda.synchronize()
                h = _lib.Hyper(opt=_lib.PDM_OPT_ADAM, lr=1e-2, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, step=t)
                comm.allreduce_optim(a[:n], p[:n], m[:n], v[:n], h, algo=algo)
                torch.cuda.synchronize()
                g = torch.zeros(n)
                for r in range(world):
                    g = g + _rand(7000 + 10 * t + r, n)
                ref_p.grad = (g * (1.0 / world)).cuda()
                ropt.step()
            torch.testing.assert_close(p[:n], ref_p.detach(), rtol=1e-5, atol=1e-6)
            mine = p[:n].clone()
            others = [torch.empty_like(mine) for _ in range(world)]
            dist.all_gather(others, mine)
            for o in others:
                assert torch.equal(o, mine), 'replicas diverged (n={}, algo={})'.format(n, algo)
        assert comm.status() == 0
        comm.close()
    finally:
        dist.destroy_process_group()


def test_comm_broadcast_allreduce_fused_optimizer(free_port):
    mp.spawn(_comm_worker, args=(_world(), free_port), nprocs=_world())


def _net_worker(rank, world, port):
    import torch.nn.functional as F
    from pytorch_distributed_mnist_b200 import Net, FusedAdam, DistributedDataParallel
    from pytorch_distributed_mnist_b200.ops.fused_step import NetStepper
    _init(rank, world, port)
    try:
        dev = torch.device('cuda', rank)
        torch.manual_seed(50 + rank)
        model = DistributedDataParallel(Net().to(dev), device_ids=[rank])
        assert model.uses_peer
        opt = FusedAdam(model.parameters(), lr=1e-3)
        st = NetStepper(dev, comm=model.comm)
        # oracle: stock DDP + Adam from the same (broadcast) init
        ref = torch.nn.Linear(784, 10).to(dev)
        ref.load_state_dict(model.module.fc.state_dict())
        ref = torch.nn.parallel.DistributedDataParallel(ref, device_ids=[rank])
        ropt = torch.optim.Adam(ref.parameters(), lr=1e-3)
        for t in range(1, 26):
            B = 64 if t % 5 else 24
            g = torch.Generator().manual_seed(1000 * t + rank)
            x = torch.randint(0, 256, (B, 784), generator=g, dtype=torch.uint8).to(dev)
            y = torch.randint(0, 10, (B,), generator=g).to(dev)
            st.train_step(x, y, opt.flat.params, opt.flat.exp_avg, opt.flat.exp_avg_sq, opt.hyper(t))
            loss = F.cross_entropy(ref(x.float().div(255).sub(MEAN).div(STD)), y)
            ropt.zero_grad()
            loss.backward()
            ropt.step()
        torch.cuda.synchronize()
        torch.testing.assert_close(model.module.fc.weight.detach(), ref.module.weight.detach(), rtol=1e-3, atol=2e-4)
        mine = opt.flat.params.clone()
        others = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(others, mine)
        for o in others:
            assert torch.equal(o, mine), 'replicas diverged in the fused LL step'
        model.comm.check_status()
        model.close()
    finally:
        dist.destroy_process_group()


def test_fused_net_step_ll_allreduce(free_port):
    mp.spawn(_net_worker, args=(_world(), free_port), nprocs=_world())


def _assert_close_after_adam(a, b, lr, steps):
    """Adam's first steps move every element by ~lr whatever the gradient's size, so tf32 / atomic-order
    noise in near-zero gradients (a hidden unit at the ReLU boundary, a dead channel) is amplified to a
    visible fraction of lr for a handful of elements.  Compare the bulk tightly and bound the outliers:
    measured on B200, 4 steps: mean drift ~1e-7, 4 of 1.18 M elements above 2e-4 (max 3.8e-4) at 8 GPUs."""
    d = (a.detach() - b.detach()).abs()
    # bounds a few times what was measured at 8 GPUs (the worst case); run with these bounds at 2 GPUs
    # (profiles/raw/r2/pytest_gpu_2gpus_native_reducer.log): a reducer or slicing bug that touches even a
    # hundred elements fails here
    assert d.max().item() <= 1e-3, 'max drift {:.3e}'.format(d.max().item())
    assert d.mean().item() < 2e-6, 'mean drift {:.3e}'.format(d.mean().item())
    outliers = int((d > 2e-4 + 2e-3 * b.detach().abs()).sum().item())
    assert outliers <= 64, '{} of {} elements drifted'.format(outliers, d.numel())


def _tc_worker(rank, world, port):
    """The persistent tensor-core kernel across GPUs: one launch per rank runs 24 steps with the LL
    exchange inside; oracle = stock DDP + torch.optim.Adam on the same per-rank batches."""
    import torch.nn.functional as F
    from pytorch_distributed_mnist_b200 import Net, FusedAdam, DistributedDataParallel
    from pytorch_distributed_mnist_b200.ops import net_tc
    from pytorch_distributed_mnist_b200.ops.engine import NetEngine
    os.environ['PDM_NET_TC_MAX_WORLD'] = '8'          # whatever the policy default: this test is about the kernel
    _init(rank, world, port)
    try:
        dev = torch.device('cuda', rank)
        if not net_tc.supported(dev):
            return
        torch.manual_seed(70 + rank)
        model = DistributedDataParallel(Net().to(dev), device_ids=[rank])
        assert model.uses_peer
        opt = FusedAdam(model.parameters(), lr=1e-3)
        eng = NetEngine(opt.flat, opt, comm=model.comm, device=dev, max_batch=256, slots=4, workers=2)
        ref = torch.nn.Linear(784, 10).to(dev)
        ref.load_state_dict(model.module.fc.state_dict())
        ref = torch.nn.parallel.DistributedDataParallel(ref, device_ids=[rank])
        ropt = torch.optim.Adam(ref.parameters(), lr=1e-3)
        B, nb, steps = 256, 6, 24
        g = torch.Generator().manual_seed(500 + rank)
        x = torch.randint(0, 256, (nb * B, 784), generator=g, dtype=torch.uint8)
        x[torch.rand(nb * B, 784, generator=g) < 0.6] = 0
        y = torch.randint(0, 10, (nb * B,), generator=g, dtype=torch.uint8)
        xd, yd = x.to(dev), y.to(dev)
        before = eng.launches
        eng.train_device(xd, yd, B, steps)                  # ONE launch: 24 steps, all-reduce inside
        assert eng.launches - before == 1
        for i in range(steps):
            s = (i % nb) * B
            loss = F.cross_entropy(ref(xd[s:s + B].float().div(255).sub(MEAN).div(STD)), yd[s:s + B].long())
            ropt.zero_grad()
            loss.backward()
            ropt.step()
        torch.cuda.synchronize()
        torch.testing.assert_close(model.module.fc.weight.detach(), ref.module.weight.detach(), rtol=1e-3, atol=2e-4)
        torch.testing.assert_close(model.module.fc.bias.detach(), ref.module.bias.detach(), rtol=1e-3, atol=2e-4)
        # a gather pass from pinned host memory continues the same run (sequence numbers carry on)
        eng.set_dataset(x, y)
        eng.place_dataset('zerocopy')
        idx = torch.randperm(nb * B, generator=torch.Generator().manual_seed(3 + rank))[:1000]
        st, _ = eng.train(idx, B)
        assert st.shape == (4, 2) and opt.step_count == steps + 4
        mine = opt.flat.params.clone()
        others = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(others, mine)
        for o in others:
            assert torch.equal(o, mine), 'replicas diverged in the persistent tensor-core kernel'
        model.comm.check_status()
        eng.close()
        model.close()
    finally:
        dist.destroy_process_group()


def test_persistent_tc_kernel_ll_allreduce(free_port):
    mp.spawn(_tc_worker, args=(_world(), free_port), nprocs=_world())


def _cnn_worker(rank, world, port, native=False, nvls=False):
    import pytorch_distributed_mnist_b200 as pdm
    if nvls:
        os.environ['PDM_NVLS'] = '1'          # reduce every bucket in the switch (normally: >= 4 GPUs, >= 8 MB)
    _init(rank, world, port)
    try:
        dev = torch.device('cuda', rank)
        torch.manual_seed(7 + rank)
        ours = pdm.DistributedDataParallel(pdm.CNN().to(dev), device_ids=[rank])
        assert ours.uses_peer and len(ours.buckets) >= 2
        assert (ours.nvls is not None) == bool(nvls)
        torch.manual_seed(7)
        stock_mod = pdm.CNN().to(dev)
        stock = torch.nn.parallel.DistributedDataParallel(stock_mod, device_ids=[rank])
        for a, b in zip(ours.module.parameters(), stock_mod.parameters()):
            assert torch.equal(a, b), 'peer broadcast from rank 0 failed'
        o_opt = pdm.FusedAdam(ours.parameters(), lr=1e-3)
        s_opt = torch.optim.Adam(stock.parameters(), lr=1e-3)
        ours.eval(); stock.eval()
        eng = None
        if native:
            # the native step executor issues the same bucket kernels from C++ on its side stream
            from pytorch_distributed_mnist_b200.ops.cnn_engine import CnnEngine
            eng = CnnEngine(ours.module, ours.flat, o_opt, ddp=ours, max_batch=32)
            assert eng.comm is not None and len(eng._ranges) >= 2
        g = torch.Generator().manual_seed(70 + rank)
        for _ in range(4):
            x = torch.randn(32, 1, 28, 28, generator=g).to(dev)
            y = torch.randint(0, 10, (32,), generator=g).to(dev)
            if eng is not None:
                eng.train_step(x, y, seed=1)
            for m, o in ((stock, s_opt),) if eng is not None else ((ours, o_opt), (stock, s_opt)):
                loss = torch.nn.functional.cross_entropy(m(x), y)
                o.zero_grad()
                loss.backward()
                o.step()
        torch.cuda.synchronize()
        for a, b in zip(ours.module.parameters(), stock_mod.parameters()):
            _assert_close_after_adam(a, b, lr=1e-3, steps=4)
        mine = ours.flat.params.clone()
        others = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(others, mine)
        for o in others:
            assert torch.equal(o, mine), 'replicas diverged in the bucketed fused path'
        assert o_opt.step_count == 4
        if eng is None:
            # the autograd path: csrc/reducer.cu launched every bucket kernel, in index order
            st = ours._reducer.state()
            assert ours._reducer.native_comm and st['launches'] == 4 * len(ours.buckets)
            assert st['order'] == list(range(len(ours.buckets))) and st['next'] == len(ours.buckets)
        # the fc bucket (4.7 MB) takes the two-shot kernel: Adam's moments are updated slice-wise by their
        # owners and must be gathered before a checkpoint; afterwards every rank holds the same complete state
        assert ours.flat.state_stale and len(ours.flat.sharded) >= 1
        ours.consolidate_optimizer_state()
        assert not ours.flat.state_stale
        for buf in (ours.flat.exp_avg, ours.flat.exp_avg_sq):
            mine = buf.clone()
            others = [torch.empty_like(mine) for _ in range(world)]
            dist.all_gather(others, mine)
            for o in others:
                assert torch.equal(o, mine), 'optimizer state differs across ranks after consolidation'
        lo, hi = sorted(ours.flat.sharded)[0]
        f_lo, f_hi = ours.flat.slice_of(lo, hi, world, (rank + 1) % world)       # a slice this rank does not own
        assert (ours.flat.exp_avg_sq[f_lo:f_hi] > 0).float().mean().item() > 0.1, 'moments of foreign slices are still empty'
        ours.comm.check_status()
        if eng is not None:
            eng.close()
        ours.close()
    finally:
        dist.destroy_process_group()


def test_bucketed_fused_allreduce_optimizer_cnn(free_port):
    mp.spawn(_cnn_worker, args=(_world(), free_port), nprocs=_world())


def test_native_cnn_step_multi_gpu(free_port):
    """csrc/cnn_step.cu driving the per-bucket fused all-reduce + Adam: same parameters as stock
    DDP + torch.optim.Adam after 4 steps, bit-identical replicas."""
    mp.spawn(_cnn_worker, args=(_world(), free_port, True), nprocs=_world())


def _nvls_ok():
    from pytorch_distributed_mnist_b200.parallel.nvls import NvlsRegion
    return NvlsRegion.supported(0)


@pytest.mark.parametrize('native', [False, True])
def test_cnn_reducer_in_the_switch(free_port, native):
    """The same two CNN checks with the flat buffers in the multicast region and every bucket reduced by
    the multimem kernel (autograd hooks path, then the native step executor)."""
    if not _nvls_ok():
        pytest.skip('no multicast (NVLS) support on this box')
    mp.spawn(_cnn_worker, args=(_world(), free_port, native, True), nprocs=_world())


def _run_worker(rank, world, port, out_dir):
    from pytorch_distributed_mnist_b200.cli import build_parser
    from pytorch_distributed_mnist_b200 import runner
    args = build_parser().parse_args(['--world-size', str(world), '--synthetic', '--train-size', str(4096 * world),
                                      '--test-size', '2048', '--batch-size', str(64 * world), '--epochs', '2', '-i',
                                      'tcp://127.0.0.1:{}'.format(port), '--checkpoint-dir', out_dir])
    args.rank = rank
    out = runner.run(args)
    torch.save(out['history'], os.path.join(out_dir, 'hist{}.pt'.format(rank)))


def test_full_job_spawn_nccl_peer(tmp_path, free_port):
    """BASELINE config 2 in miniature: Net, batch 64/GPU, mp.spawn-style ranks, peer path."""
    world = _world()
    mp.spawn(_run_worker, args=(world, free_port, str(tmp_path)), nprocs=world)
    hists = [torch.load(str(tmp_path / 'hist{}.pt'.format(r))) for r in range(world)]
    for h in hists:
        assert h[1]['test_acc'] > 0.85
        # every rank evaluates the full test set with identical parameters
        assert h[1]['test_acc'] == hists[0][1]['test_acc'] and abs(h[1]['test_loss'] - hists[0][1]['test_loss']) < 1e-9
    assert os.path.isfile(str(tmp_path / 'checkpoint_1.pth.tar'))


def _timeout_worker(rank, world, port):
    import time
    from pytorch_distributed_mnist_b200.parallel.peer import PeerComm
    _init(rank, world, port)
    try:
        dev = torch.device('cuda', rank)
        comm = PeerComm(rank, world, dev, data_bytes=4096, timeout_s=1.0)
        comm.barrier()                                   # everyone: a good collective first
        torch.cuda.synchronize()
        assert comm.status() == 0
        if rank == 0:
            t0 = time.time()
            comm.barrier()                               # the peers never join this one
            torch.cuda.synchronize()                     # returns: the wait is bounded, the GPU is not hung
            assert time.time() - t0 < 20
            st = comm.status()
            assert st != 0 and 1 <= st - 1 < world, st   # 1 + a rank that did not show up
            with pytest.raises(RuntimeError, match='did not reach a collective'):
                comm.check_status()
        else:
            time.sleep(4.0)                              # "dead" peer: skips the collective
        dist.barrier()
        comm.close()
    finally:
        dist.destroy_process_group()


def test_dead_peer_is_reported_not_hung(free_port):
    """a rank that never reaches a collective must produce a diagnosable error
    on its peers (bounded device-side wait + status word), not a hung GPU."""
    mp.spawn(_timeout_worker, args=(2, free_port), nprocs=2)


def _nvls_worker(rank, world, port):
    from pytorch_distributed_mnist_b200.ops import _lib
    from pytorch_distributed_mnist_b200.parallel.nvls import NvlsRegion
    _init(rank, world, port)
    try:
        dev = torch.device('cuda', rank)
        n = 1 << 18
        nv = NvlsRegion(rank, world, dev, data_bytes=n * 4 * 5 + 8192)
        g, out, p, m, v = (nv.data_tensor(n) for _ in range(5))
        g.copy_(_rand(300 + rank, n))
        torch.cuda.synchronize()
        dist.barrier()
        nv.allreduce(g, out, average=True)
        torch.cuda.synchronize()
        want = sum(_rand(300 + r, n) for r in range(world)) / world
        torch.testing.assert_close(out.cpu(), want, rtol=1e-5, atol=1e-6)
        # fused Adam: every replica must end with the same parameters as torch.optim.Adam on the mean gradient
        p.copy_(_rand(7, n))
        ref_p = _rand(7, n).to(dev).requires_grad_(True)
        ref_p.grad = want.to(dev)
        torch.optim.Adam([ref_p], lr=1e-3).step()
        h = _lib.Hyper(opt=_lib.PDM_OPT_ADAM, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, step=1)
        torch.cuda.synchronize()
        dist.barrier()
        nv.allreduce_optim(g, p, m, v, h)
        torch.cuda.synchronize()
        torch.testing.assert_close(p, ref_p.detach(), rtol=1e-5, atol=1e-6)
        nv.check_status()
        nv.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(os.environ.get('PDM_TEST_NVLS', '') in ('', '0'),
                    reason='experimental NVLS path: opt in with PDM_TEST_NVLS=1 (not yet validated on hardware)')
def test_nvls_allreduce_optim(free_port):
    mp.spawn(_nvls_worker, args=(_world(), free_port), nprocs=_world())
