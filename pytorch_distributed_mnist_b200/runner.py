"""``run(args)`` -- the per-process job (ref:163-255), device-agnostic.

Order of operations follows SURVEY 3.3-3.6: init the process group from the
explicit ``tcp://`` URL + ``world_size`` + ``rank`` (ref:167-168); divide the
node-total ``--batch-size`` (floor) and ``--workers`` (ceil) by the number of
participants (ref:174-175); ``gpu_id = rank`` (ref:180); build the model, wrap
it, build the optimizer, optionally resume (ref:196-214), build both loaders
(even for ``--evaluate``, ref:218-221), then evaluate-only or the epoch loop
with rank-0 checkpointing (ref:225-255).  Print formats are the reference's
(SURVEY 5.5).

Deliberate differences, each because the reference cannot do it at all:

* runs without GPUs (``--device cpu`` / gloo): the divisor is then
  ``world_size`` rather than ``torch.cuda.device_count()``;
* ``torch.cuda.set_device`` is called (the reference never does; our kernels
  and IPC mapping need the current device to be the rank's);
* the process group is destroyed on the way out, after a barrier, so rank 0's
  last checkpoint write cannot race a peer's teardown.
"""
from __future__ import annotations

import os

import torch
import torch.backends.cudnn as cudnn
import torch.distributed as dist

from . import checkpoint as ckpt
from .data.loader import MNISTDataLoader, distributed_is_initialized
from .models import build_model
from .optim import build_optimizer
from .parallel.ddp import DistributedDataParallel
from .schedule import adjust_learning_rate
from .trainer import Trainer

best_acc = 0


def pick_device(args, rank):
    want = getattr(args, 'device', 'auto')
    if want == 'cpu' or (want == 'auto' and (not torch.cuda.is_available() or args.backend == 'gloo' and
                                             torch.cuda.device_count() <= rank)):
        return torch.device('cpu')
    if not torch.cuda.is_available():
        raise RuntimeError('--device cuda requested but no CUDA device is visible')
    if rank >= torch.cuda.device_count():
        raise RuntimeError('rank {} has no GPU: only {} visible (the reference requires world_size == GPU count, '
                           'ref:350-351)'.format(rank, torch.cuda.device_count()))
    return torch.device('cuda', rank)


def run(args):
    global best_acc

    # initialize the process group
    dist.init_process_group(backend=args.backend, init_method=args.init_method,
                            world_size=args.world_size, rank=args.rank)
    try:
        return _run(args)
    finally:
        if distributed_is_initialized():
            try:
                dist.barrier()
            except Exception:  # noqa: BLE001 - a dead peer must not mask the real error
                pass
            dist.destroy_process_group()


def _run(args):
    global best_acc

    rank = args.rank
    device = pick_device(args, rank)
    ngpus = torch.cuda.device_count()
    # --batch-size and --workers are node totals (ref:174-175).  The reference divides by the GPU count and
    # asserts world_size == GPU count (ref:350-351); dividing by world_size is the same thing whenever that
    # holds and keeps the global batch at --batch-size when fewer ranks than visible GPUs are started.
    divisor = args.world_size
    args.batch_size = int(args.batch_size / divisor)
    args.workers = int((args.workers + divisor - 1) / divisor)
    if device.type == 'cuda':
        torch.cuda.set_device(device)

    print("rank: {}, device count: {}, workers:{}".format(rank, ngpus, args.workers))
    if getattr(args, 'print_topology', False) and rank == 0:
        from .parallel import topology
        print(topology.discover().summary(range(args.world_size) if device.type == 'cuda' else None))

    if getattr(args, 'precision', 'tf32') != 'tf32':
        from .ops import gemm as _gemm
        _gemm.set_precision(args.precision)
    model = build_model(getattr(args, 'arch', 'linear'), getattr(args, 'width', 1))
    if distributed_is_initialized():
        model.to(device)
        store = None
        if getattr(args, 'store', 'torch') == 'native' and device.type == 'cuda' and args.world_size > 1:
            from .parallel.store import store_for_job
            store = store_for_job(args.init_method, rank)
        model = DistributedDataParallel(
            model, device_ids=[device.index] if device.type == 'cuda' else None,
            comm=getattr(args, 'comm', 'auto'), store=store,
            comm_timeout_s=float(getattr(args, 'comm_timeout', 30.0)))

    optimizer = build_optimizer(getattr(args, 'optimizer', 'adam'), model.parameters(), args)

    # optionally resume from a checkpoint ('auto' = the newest one in the checkpoint directory)
    if args.resume == 'auto':
        found = ckpt.latest_checkpoint(getattr(args, 'checkpoint_dir', None) or ckpt.CHECKPOINT_DIR)
        args.resume = found if found is not None else 'auto'
    if args.resume:
        if os.path.isfile(args.resume):
            print("=> loading checkpoint '{}'".format(args.resume))
            # Map model to be loaded to current device.
            checkpoint = ckpt.load_checkpoint(args.resume, map_location=device)
            args.start_epoch = checkpoint['epoch']
            best_acc = checkpoint['best_acc']
            print("best_acc: {}".format(best_acc))
            model.load_state_dict(ckpt.adapt_state_dict_keys(checkpoint['state_dict'], model))
            optimizer.load_state_dict(checkpoint['optimizer'])
            print("=> loaded checkpoint '{}' (epoch {})".format(args.resume, checkpoint['epoch']))
        else:
            print("=> no checkpoint found at '{}'".format(args.resume))

    cudnn.benchmark = True

    data_kw = dict(synthetic=True if getattr(args, 'synthetic', False) else None,
                   seed=getattr(args, 'data_seed', 1234))
    train_loader = MNISTDataLoader(args.root, args.batch_size, num_workers=args.workers, train=True,
                                   size=getattr(args, 'train_size', None), **data_kw)
    test_loader = MNISTDataLoader(args.root, args.batch_size, num_workers=args.workers, train=False,
                                  size=getattr(args, 'test_size', None), **data_kw)

    trainer = Trainer(model, optimizer, train_loader, test_loader, device,
                      fused=getattr(args, 'fused', 'auto'),
                      data_placement=getattr(args, 'data_placement', 'auto'))
    chk_dir = getattr(args, 'checkpoint_dir', None) or ckpt.CHECKPOINT_DIR
    try:
        if args.evaluate:
            test_loss, test_acc = trainer.evaluate()
            print('test loss: {}, test acc: {}.'.format(test_loss, test_acc))
            return {'test_loss': test_loss.average, 'test_acc': test_acc.accuracy}

        history = []
        for epoch in range(args.start_epoch, args.epochs):
            train_loader.set_sample_epoch(epoch)
            adjust_learning_rate(optimizer, epoch, args)

            # train for one epoch
            train_loss, train_acc = trainer.train()
            test_loss, test_acc = trainer.evaluate()

            print(
                'Epoch: {}/{},'.format(epoch, args.epochs),
                'train loss: {}, train acc: {},'.format(train_loss, train_acc),
                'test loss: {}, test acc: {}.'.format(test_loss, test_acc)
            )
            history.append({'epoch': epoch, 'train_loss': train_loss.average, 'train_acc': train_acc.accuracy,
                            'test_loss': test_loss.average, 'test_acc': test_acc.accuracy})

            # remember best acc and save checkpoint
            is_best = test_acc.accuracy > best_acc
            best_acc = max(test_acc.accuracy, best_acc)

            # every rank: the two-shot fused update keeps optimizer-state slices on their owners
            if hasattr(model, 'consolidate_optimizer_state'):
                model.consolidate_optimizer_state()
            # only save checkpoints in the first gpu
            if args.rank == 0:
                ckpt.save_checkpoint(ckpt.make_state(epoch, model, best_acc, optimizer), is_best, epoch, chk_dir)
        return {'history': history, 'best_acc': best_acc}
    finally:
        trainer.close()
        if hasattr(model, 'close'):
            if device.type == 'cuda':
                torch.cuda.synchronize(device)
            model.close()
