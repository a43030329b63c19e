#!/usr/bin/env python
"""bench.py -- headline benchmark: training throughput (images/s) of the reference model
(``Net`` = Linear(784, 10), Adam lr 1e-3, fp32, batch 512 per GPU -- BASELINE.json config 3,
weak scaling) on N GPUs of one node, synthetic MNIST-shaped data, random-init weights.

    python bench.py --gpus 1 --steps K --warmup W
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # the unmodified reference from baseline/_ref

Two numbers per run, both whole-job aggregates (max over ranks of the time):

* ``value``  -- device-timed: exactly K training steps (forward, loss, backward, gradient
  all-reduce, optimizer update, metric accumulation) on inputs already resident on the
  device, cycling through a pool of distinct batches larger than L2, CUDA events on the
  launching stream.
* ``e2e``    -- the same metric through the public API a user calls (``Trainer.train()``
  over an ``MNISTDataLoader``): every step includes the host->device copy of that step's
  inputs from pinned host memory and a device->host read of the step's {loss, correct}.
  ``e2e.value`` is images / WALL time of the ``Trainer.train()`` call (host clock around the
  call, barrier + synchronize on both sides, max over ranks); the CUDA-event span of the same
  pass is reported beside it as ``device_ms_per_step``.

The JSON line also carries, measured in separate processes (one per rank, same launch geometry)
after the headline numbers are final so that they can neither disturb nor block them:

* ``cnn_extension`` -- the training step of the CNN that BASELINE.json's configs 2-4 name (not in
  the reference): our native step executor vs stock DistributedDataParallel on cuDNN/cuBLAS TF32 +
  ``torch.optim.Adam``, width 1 at batch 512/GPU and width 8 at batch 128/GPU, at N GPUs
  (``benchmarks/bench_cnn.py``; ``--no-cnn`` skips it);
* ``allreduce_us`` (N > 1) -- the other half of BASELINE.json's metric: the gradient all-reduce
  per step for the 31,400-byte ``Net`` bucket and for 64 MB, ours (fused with 1/W and Adam) vs
  NCCL (+ Adam) (``benchmarks/allreduce_sweep.py``; ``--no-allreduce`` skips it).

The reference arm builds the reference's own ``Net`` / ``MNISTDataLoader`` / ``Trainer``
exactly as its ``run()`` does (ref:185-223: stock DDP, ``torch.optim.Adam``) and times its
``Trainer.train()``; nothing from this package is on its timed path (this package only
writes the synthetic idx files the reference's torchvision loader reads, beforehand).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

PER_GPU_BATCH = 512
POOL_BATCHES = 512          # 512 x 512 x 785 B = 206 MB of uint8 inputs  (> 126 MB L2)
REF_POOL_BATCHES = 128      # 128 x 512 x 3,144 B = 206 MB of fp32/int64 inputs
REF_E2E_MAX_STEPS = 100     # the reference's PIL loader decodes ~25k images/s per worker
# BASELINE.json's metric string.  Its "MNIST-CNN" is the reference's `Net`, which is a single
# Linear(784, 10): that model, through the reference's own code, is what both arms run.
METRIC = 'MNIST-CNN train images/sec (device-timed, max over ranks)'


# ------------------------------------------------------------------------------ clocks
class ClockSampler(object):
    """Samples SM clock + throttle reasons of one GPU through NVML while a region runs."""
    REASONS = {0x8: 'hw_slowdown', 0x40: 'hw_thermal_slowdown', 0x20: 'sw_thermal_slowdown',
               0x4: 'sw_power_cap', 0x80: 'hw_power_brake_slowdown', 0x2: 'applications_clocks_setting'}

    def __init__(self, index, period_s=0.01):
        self.index, self.period = index, period_s
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:  # noqa: BLE001
            self.nv = None

    def _once(self):
        try:
            self.samples.append(float(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
            mask = int(self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
            for bit, name in self.REASONS.items():
                if mask & bit:
                    self.reasons.add(name)
        except Exception:  # noqa: BLE001
            pass

    def _loop(self):
        while not self._stop.is_set():
            self._once()
            self._stop.wait(self.period)

    def __enter__(self):
        if self.nv is not None:
            self._once()
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *exc):
        if self._thread is not None:
            self._once()
            self._stop.set()
            self._thread.join()

    def summary(self):
        if not self.samples:
            return {'sm_mhz': None, 'sm_max_mhz': self.max_mhz, 'reasons': ['unavailable'], 'samples': 0}
        s = sorted(self.samples)
        return {'sm_mhz': s[len(s) // 2], 'sm_max_mhz': self.max_mhz, 'reasons': sorted(self.reasons),
                'samples': len(s)}


# ------------------------------------------------------------------------------ distributed helpers
def dist_env():
    rank = int(os.environ.get('RANK', 0))
    local = int(os.environ.get('LOCAL_RANK', rank))
    world = int(os.environ.get('WORLD_SIZE', 1))
    return rank, local, world


def init_dist(world, rank, local, port_hint=29533):
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    if world > 1 or 'MASTER_ADDR' in os.environ:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        os.environ.setdefault('MASTER_PORT', str(port_hint))
        dist.init_process_group('nccl', init_method='env://', world_size=world, rank=rank,
                                device_id=torch.device('cuda', local))
    else:
        import socket
        s = socket.socket()
        s.bind(('127.0.0.1', 0))
        port = s.getsockname()[1]
        s.close()
        dist.init_process_group('nccl', init_method='tcp://127.0.0.1:{}'.format(port), world_size=1, rank=0,
                                device_id=torch.device('cuda', local))
    return dist


def sync_all(dist, device):
    import torch
    torch.cuda.synchronize(device)
    if dist.get_world_size() > 1:
        dist.barrier()
    torch.cuda.synchronize(device)


def max_over_ranks(dist, value, device):
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    if dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


# ------------------------------------------------------------------------------ our arm
def run_ours(opts):
    import torch
    import pytorch_distributed_mnist_b200 as pdm
    rank, local, world = dist_env()
    if world != opts.gpus:
        raise SystemExit('bench.py: --gpus {} but WORLD_SIZE={} (launch N>1 with torch.distributed.run)'
                         .format(opts.gpus, world))
    device = torch.device('cuda', local)
    dist = init_dist(world, rank, local)
    B, K, W = PER_GPU_BATCH, opts.steps, opts.warmup
    torch.manual_seed(0)
    model = pdm.DistributedDataParallel(pdm.Net().to(device), device_ids=[local], comm=opts.comm)
    optimizer = pdm.FusedAdam(model.parameters(), lr=1e-3)

    # loaders of exactly W and K steps per rank (the distributed sampler shards N x steps x B samples)
    warm_loader = pdm.MNISTDataLoader('unused', B, num_workers=opts.workers, train=True, synthetic=True,
                                      size=max(1, W) * B * world)
    # the timed pass walks the standard 60,000-sample set (the reference's working set, 47 MB) for
    # exactly K steps per rank: successive epoch permutations, sharded by the distributed sampler
    timed_loader = pdm.MNISTDataLoader('unused', B, num_workers=opts.workers, train=True, synthetic=True,
                                       seed=4321, samples_per_epoch=K * B)
    trainer = pdm.Trainer(model, optimizer, warm_loader, warm_loader, device, fused='on', ctas=opts.ctas,
                          loader_workers=opts.workers, data_placement=opts.placement)
    engine = trainer.engine

    # device-resident pool for the kernel-only number
    g = torch.Generator(device=device).manual_seed(1 + rank)
    pool_x = torch.randint(0, 256, (POOL_BATCHES * B, 784), dtype=torch.uint8, device=device, generator=g)
    pool_y = torch.randint(0, 10, (POOL_BATCHES * B,), dtype=torch.uint8, device=device, generator=g)

    with ClockSampler(local) as clocks:
        # ---- value: K steps, inputs resident (one launch for the whole pass when the persistent kernel takes it) ----
        sync_all(dist, device)
        engine.train_device(pool_x, pool_y, B, max(W, 3), ctas=opts.ctas)
        sync_all(dist, device)
        launches0 = engine.launches
        t0 = time.perf_counter()
        ms = engine.train_device(pool_x, pool_y, B, K, ctas=opts.ctas)
        sync_all(dist, device)
        wall_ms = (time.perf_counter() - t0) * 1e3
        launches = engine.launches - launches0
        ms = max_over_ranks(dist, ms, device)
        wall_ms = max_over_ranks(dist, wall_ms, device)

        # ---- the same K steps on fp32 inputs (what the reference arm reads: 3,136 B/sample + int64 labels) ----
        del pool_x, pool_y
        nb32 = REF_POOL_BATCHES
        pool_xf = torch.randn(nb32 * B, 784, device=device, generator=g)
        pool_yl = torch.randint(0, 10, (nb32 * B,), device=device, generator=g)
        sync_all(dist, device)
        engine.train_device(pool_xf, pool_yl, B, max(W, 3), ctas=opts.ctas)
        sync_all(dist, device)
        ms_f32 = max_over_ranks(dist, engine.train_device(pool_xf, pool_yl, B, K, ctas=opts.ctas), device)
        del pool_xf, pool_yl

        # ---- e2e: Trainer.train() over a host dataset ----
        trainer.train_loader = warm_loader
        warm_loader.set_sample_epoch(0)
        trainer.train()
        trainer.train_loader = timed_loader          # binds (page-locks / hands over) the dataset: set-up, not a step
        timed_loader.set_sample_epoch(0)             # ref:231 -- the caller's per-epoch call, before train()
        sync_all(dist, device)
        t0 = time.perf_counter()
        loss_avg, acc = trainer.train()
        torch.cuda.synchronize(device)               # train() has already read every step's result back
        e2e_wall_ms = (time.perf_counter() - t0) * 1e3
        sync_all(dist, device)
        e2e_wall_ms = max_over_ranks(dist, e2e_wall_ms, device)
        e2e_dev_ms = max_over_ranks(dist, trainer.last_train_ms, device)
    clk = clocks.summary()

    images = K * B * world
    value = images / (ms / 1e3)
    e2e_value = images / (e2e_wall_ms / 1e3)
    out = {
        'metric': METRIC, 'value': value, 'unit': 'images/s', 'n_gpus': world, 'steps': K, 'warmup': max(W, 3),
        'ms_per_step': ms / K, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
        'dtype': 'fp32', 'data': 'synthetic (uint8 MNIST-shaped, random-init weights)',
        'config': {'model': 'Net = Linear(784,10) (reference model, ref:119-126)', 'optimizer': 'Adam lr=1e-3',
                   'global_batch': B * world, 'per_gpu_batch': B, 'seq_len': None,
                   'parallelism': 'dp{}'.format(world),
                   'comm': 'peer-memory LL all-reduce fused in the step kernel' if getattr(model, 'uses_peer', False) and world > 1
                           else ('single GPU' if world == 1 else 'torch.distributed fallback'),
                   'l2_policy': 'inputs larger than L2: {} distinct batches = {:.0f} MB cycled'.format(
                       POOL_BATCHES, POOL_BATCHES * B * 785 / 1e6),
                   'inputs': 'raw uint8 pixels, normalised in-kernel (bit-exact ToTensor+Normalize); fp32 math',
                   'kernel': ('persistent tcgen05 cluster kernel, one launch for the K steps (csrc/net_tc.cu; bf16 operand '
                              'terms with exact products, fp32 accumulation in TMEM)' if launches == 1 and K > 1
                              else 'per-step fused FMA kernel, one launch per step (csrc/net_step.cu)')},
        'wall_ms_per_step': wall_ms / K,
        'fp32_inputs': {'value': images / (ms_f32 / 1e3), 'ms_per_step': ms_f32 / K,
                        'note': 'the per-step FMA kernel (csrc/net_step.cu) on already-normalised fp32 pixels + int64 labels, the bytes '
                                'the reference arm reads (3,144 B/sample); {} distinct batches = {:.0f} MB cycled'.format(
                                    nb32, nb32 * B * 3144 / 1e6)},
        'clocks': clk,
        'e2e': {'value': e2e_value, 'unit': 'images/s',
                'h2d_bytes_per_step': B * 785 + (B * 4 if trainer.data_placement == 'zerocopy' else 0),
                'd2h_bytes_per_step': 8,
                'ms_per_step': e2e_wall_ms / K, 'timing': 'host wall clock around Trainer.train(), max over ranks',
                'device_ms_per_step': e2e_dev_ms / K, 'steps': K,
                'api': 'Trainer.train() over MNISTDataLoader',
                'input_path': {
                    'staged': 'loader threads gather each shuffled batch into a pinned slot; one cudaMemcpyAsync H2D '
                              '(B*785 bytes) per step',
                    'zerocopy': 'every step the kernel reads its shuffled rows (B*785 bytes) over PCIe straight from the '
                                'page-locked (cudaHostRegister) host dataset: a feeder cluster of the persistent kernel gathers '
                                'whole rows into a device ring one to four blocks ahead of the training cluster; index plan '
                                'uploaded as B*4 bytes/step',
                }.get(trainer.data_placement, trainer.data_placement),
                'result_path': 'kernel stores {loss, correct} (8 bytes) into mapped pinned host memory every step',
                'host_cores_per_rank': (os.cpu_count() or 0) // max(1, world),
                'final_loss': loss_avg.average},
        'gpu_launches': launches,
        'impl': 'ours',
    }
    # everything below runs in child processes (one per rank) once the headline numbers are final
    trainer.close()
    sync_all(dist, device)
    model.close()
    sync_all(dist, device)
    if opts.cnn:
        out['cnn_extension'] = cnn_extension_numbers(rank, local, world)
    if world > 1 and opts.allreduce:
        out['allreduce_us'] = allreduce_numbers(rank, local, world)
    if rank == 0:
        print(json.dumps(out), flush=True)
    sync_all(dist, device)
    dist.destroy_process_group()


# ------------------------------------------------------------------------------ CNN extension (supplementary)
def run_json_lines(cmd, timeout_s, env=None):
    """Run ``cmd`` in its own process and return the JSON objects it printed, one per line.  Never
    raises: a failure comes back as ``[{'unavailable': reason}]`` (the headline number must not depend on it)."""
    import subprocess
    try:
        proc = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout_s, env=env, cwd=ROOT)
    except subprocess.TimeoutExpired:
        return [{'unavailable': 'timed out after {} s'.format(timeout_s)}]
    except OSError as e:
        return [{'unavailable': '{}: {}'.format(type(e).__name__, e)}]
    rows = []
    for line in proc.stdout.splitlines():
        line = line.strip()
        if line.startswith('{') and line.endswith('}'):
            try:
                rows.append(json.loads(line))
            except ValueError:
                pass
    if proc.returncode != 0 or not rows:
        tail = (proc.stderr or proc.stdout or '').strip().splitlines()[-1:] or ['no output']
        rows.append({'unavailable': 'exit code {}: {}'.format(proc.returncode, tail[0][:200])})
    return rows


def child_env(rank, local, world, port):
    """Environment of this rank's child process: the same rank geometry, a rendezvous of its own.  Under
    torchrun the parent's TORCHELASTIC_* variables (in particular TORCHELASTIC_USE_AGENT_STORE, which turns
    env:// into a client of the agent's store) must not leak."""
    env = {k: v for k, v in os.environ.items()
           if k not in ('GROUP_RANK', 'LOCAL_WORLD_SIZE', 'ROLE_RANK', 'ROLE_WORLD_SIZE', 'GROUP_WORLD_SIZE', 'ROLE_NAME')
           and not k.startswith('TORCHELASTIC_')}
    env.update({'RANK': str(rank), 'LOCAL_RANK': str(local), 'WORLD_SIZE': str(world), 'MASTER_ADDR': '127.0.0.1',
                'MASTER_PORT': str(port)})
    return env


def child_port(offset):
    """A rendezvous port every rank derives identically (children of all ranks must meet on it)."""
    base = int(os.environ.get('MASTER_PORT', '29533'))
    return 20000 + (base * 7 + 131 * offset) % 20000


def cnn_extension_numbers(rank, local, world):
    """BASELINE.json's configs 2-4 also name an MNIST CNN (an extension: the reference has none).
    Its training step through our native step executor vs the same module under stock DistributedDataParallel on
    cuDNN/cuBLAS TF32 + torch.optim.Adam, at this run's GPU count: every rank starts benchmarks/bench_cnn.py as a
    child with its own rank geometry (device-timed there, max over ranks), rank 0 reports."""
    out = {}
    for k, (width, batch) in enumerate(((1, 512), (8, 128))):
        rows = run_json_lines([sys.executable, os.path.join(ROOT, 'benchmarks', 'bench_cnn.py'), '--width', str(width),
                               '--batch', str(batch), '--steps', '40', '--warmup', '10', '--impl', 'engine,stock'],
                              timeout_s=180, env=child_env(rank, local, world, child_port(1 + k)))
        entry = {'per_gpu_batch': batch, 'n_gpus': world}
        for r in rows:
            if 'unavailable' in r:
                entry['unavailable'] = r['unavailable']
            elif r.get('impl') == 'engine':
                entry.update({'ms_per_step': r['ms_per_step'], 'images_per_s': r['images_per_s'], 'params': r['params']})
            elif r.get('impl') == 'stock':
                entry.update({'stock_ms_per_step': r['ms_per_step'], 'stock_images_per_s': r['images_per_s']})
        if 'ms_per_step' in entry and 'stock_ms_per_step' in entry:
            entry['speedup_vs_stock'] = entry['stock_ms_per_step'] / entry['ms_per_step']
        out['width{}'.format(width)] = entry
    out['note'] = ('supplementary, device-timed, max over ranks, weak scaling: models.CNN (PyTorch-examples MNIST net x width) '
                   'training step on the tcgen05/TMA kernels + native step executor + peer-memory fused all-reduce/Adam vs the '
                   'same module under torch DistributedDataParallel on cuDNN/cuBLAS TF32 + torch.optim.Adam')
    return out


def allreduce_numbers(rank, local, world):
    """The gradient all-reduce per step (BASELINE.json metric, second half; config 5) for the reference model's one
    31,400-byte bucket and for 64 MB: NCCL all-reduce, NCCL + 1/W + torch.optim.Adam (what stock DDP + optimizer do per
    bucket), our peer-memory all-reduce, and ours fused with 1/W + Adam (what our reducer launches).  us per call,
    CUDA events, max over ranks; measured by benchmarks/allreduce_sweep.py in child processes."""
    rows = run_json_lines([sys.executable, os.path.join(ROOT, 'benchmarks', 'allreduce_sweep.py'), '--sizes',
                           '31408,67108864', '--iters', '100', '--nvls'],
                          timeout_s=180, env=child_env(rank, local, world, child_port(5)))
    out = {'n_gpus': world}
    for r in rows:
        if 'unavailable' in r:
            out['unavailable'] = r['unavailable']
        elif 'bytes' in r:
            out[str(r['bytes'])] = {k: v for k, v in r.items() if k not in ('bytes', 'world')}
    return out


# ------------------------------------------------------------------------------ reference arm
def run_reference(opts):
    ref_dir = os.path.join(ROOT, 'baseline', '_ref')
    try:
        if not os.path.isfile(os.path.join(ref_dir, 'multi_proc_single_gpu.py')):
            raise ImportError('baseline/_ref/multi_proc_single_gpu.py missing (run baseline/install_reference.sh)')
        sys.path.insert(0, ref_dir)
        import multi_proc_single_gpu as ref   # the unmodified reference module
    except Exception as e:  # noqa: BLE001
        print(json.dumps({'impl': 'reference', 'unavailable': '{}: {}'.format(type(e).__name__, e)}), flush=True)
        return
    import torch
    import torch.nn as nn
    rank, local, world = dist_env()
    device = torch.device('cuda', local)
    dist = init_dist(world, rank, local)
    B, K, W = PER_GPU_BATCH, opts.steps, max(opts.warmup, 3)
    e2e_steps = min(K, REF_E2E_MAX_STEPS)

    # synthetic idx files where torchvision's MNIST(download=True) looks for them (offline stand-in
    # for the download; written before anything is timed)
    from pytorch_distributed_mnist_b200.data import mnist
    tmp = os.environ.get('PDM_BENCH_TMP', tempfile.gettempdir())
    roots = {'warm': os.path.join(tmp, 'pdm_ref_warm_{}_{}'.format(world, W)),
             'timed': os.path.join(tmp, 'pdm_ref_timed_{}_{}'.format(world, e2e_steps))}
    if rank == 0:
        mnist.write_idx_tree(roots['warm'], seed=1234, train_size=min(W, 10) * B * world, test_size=B)
        mnist.write_idx_tree(roots['timed'], seed=4321, train_size=e2e_steps * B * world, test_size=B)
    if world > 1:
        dist.barrier()

    # ---- exactly what ref.run() does at ref:185-191 ----
    torch.manual_seed(0)
    model = ref.Net()
    model.cuda(device)
    model = nn.parallel.DistributedDataParallel(model, device_ids=[local])
    optimizer = torch.optim.Adam(model.parameters(), lr=1e-3)
    workers = int((4 + world - 1) / world)     # ref:175 with the default --workers 4

    # device-resident pool (fp32 normalised images + int64 labels, what the reference's loader yields)
    g = torch.Generator(device=device).manual_seed(1 + rank)
    pool = [(torch.randn(B, 1, 28, 28, device=device, generator=g),
             torch.randint(0, 10, (B,), device=device, generator=g)) for _ in range(REF_POOL_BATCHES)]

    class Cycle(object):          # an iterable of K batches; .cuda() on them is a no-op
        def __init__(self, n):
            self.n = n

        def __iter__(self):
            return (pool[i % len(pool)] for i in range(self.n))

    trainer = ref.Trainer(model, optimizer, Cycle(W), None, device)
    with ClockSampler(local) as clocks:
        sync_all(dist, device)
        trainer.train()                                        # warm-up
        trainer.train_loader = Cycle(K)
        sync_all(dist, device)
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        ev0.record()
        trainer.train()
        ev1.record()
        sync_all(dist, device)
        wall_ms = max_over_ranks(dist, (time.perf_counter() - t0) * 1e3, device)
        ms = max_over_ranks(dist, ev0.elapsed_time(ev1), device)

        # ---- e2e: the reference's own MNISTDataLoader (torchvision + PIL + DataLoader workers) ----
        warm_loader = ref.MNISTDataLoader(roots['warm'], B, num_workers=workers, train=True)
        timed_loader = ref.MNISTDataLoader(roots['timed'], B, num_workers=workers, train=True)
        trainer.train_loader = warm_loader
        trainer.train()
        trainer.train_loader = timed_loader
        assert len(timed_loader) == e2e_steps, (len(timed_loader), e2e_steps)
        sync_all(dist, device)
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        ev0.record()
        loss_avg, acc = trainer.train()
        ev1.record()
        sync_all(dist, device)
        e2e_wall_ms = max_over_ranks(dist, (time.perf_counter() - t0) * 1e3, device)
        e2e_ms = max_over_ranks(dist, ev0.elapsed_time(ev1), device)
    clk = clocks.summary()

    out = {
        'metric': METRIC, 'value': K * B * world / (ms / 1e3), 'unit': 'images/s', 'n_gpus': world, 'steps': K,
        'warmup': W, 'ms_per_step': ms / K, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
        'dtype': 'fp32', 'data': 'synthetic (MNIST-shaped, random-init weights)',
        'config': {'model': 'reference Net = Linear(784,10)', 'optimizer': 'torch.optim.Adam lr=1e-3',
                   'global_batch': B * world, 'per_gpu_batch': B, 'seq_len': None, 'parallelism': 'dp{}'.format(world),
                   'comm': 'stock DistributedDataParallel + NCCL',
                   'l2_policy': 'inputs larger than L2: {} distinct fp32 batches = {:.0f} MB cycled'.format(
                       REF_POOL_BATCHES, REF_POOL_BATCHES * B * 3144 / 1e6)},
        'wall_ms_per_step': wall_ms / K,
        'clocks': clk,
        'e2e': {'value': e2e_steps * B * world / (e2e_ms / 1e3), 'unit': 'images/s',
                'h2d_bytes_per_step': B * (3136 + 8), 'd2h_bytes_per_step': 12, 'ms_per_step': e2e_ms / e2e_steps,
                'wall_ms_per_step': e2e_wall_ms / e2e_steps, 'steps': e2e_steps,
                'api': 'reference Trainer.train() over reference MNISTDataLoader (pageable .cuda() per step)',
                'final_loss': loss_avg.average},
        'gpu_launches': None,
        'impl': 'reference',
    }
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=2000)
    ap.add_argument('--warmup', type=int, default=100)
    ap.add_argument('--impl', choices=('ours', 'reference'), default='ours')
    ap.add_argument('--workers', type=int, default=6, help='loader threads per rank (ours)')
    ap.add_argument('--ctas', type=int, default=0, help='CTAs of the fused step kernel (0 = from batch)')
    ap.add_argument('--comm', choices=('auto', 'peer', 'group'), default='auto')
    ap.add_argument('--placement', choices=('auto', 'staged', 'zerocopy'), default='auto',
                    help='e2e input path (ours); both move every step\'s inputs from pinned host memory')
    ap.add_argument('--no-cnn', dest='cnn', action='store_false',
                    help='skip the supplementary CNN-extension numbers (ours)')
    ap.add_argument('--no-allreduce', dest='allreduce', action='store_false',
                    help='skip the all-reduce latency numbers (ours, --gpus > 1)')
    opts = ap.parse_args()
    if opts.steps < 1:
        raise SystemExit('--steps must be >= 1')
    if opts.impl == 'reference':
        run_reference(opts)
    else:
        run_ours(opts)


if __name__ == '__main__':
    main()
