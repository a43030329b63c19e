"""Command line: the reference's 15 flags verbatim (ref:289-334)
plus the extensions SURVEY 7.2 step 1 calls for.

Launch-mode selection needs no source edit: the reference makes
users comment/uncomment ref:355/359; here ``--launcher auto`` picks *launch*
mode when a launcher's environment (``LOCAL_RANK`` / ``RANK``) or an explicit
``--local-rank`` is present, else *spawn* mode.  Both ``--local-rank`` (what
``torch.distributed.launch`` passes since torch 2.0) and ``--local_rank``
(what the reference declares, ref:321) are accepted.
"""
from __future__ import annotations

import argparse
import os
import random
import sys
import warnings

import torch
import torch.backends.cudnn as cudnn


def build_parser():
    parser = argparse.ArgumentParser(description='B200-native distributed MNIST trainer')
    # ---- the reference surface, unchanged (ref:289-334) ----
    parser.add_argument('--root', type=str, default='data')
    parser.add_argument('-j', '--workers', default=4, type=int, metavar='N',
                        help='number of data loading workers (default: 4)')
    parser.add_argument('--epochs', type=int, default=20)
    parser.add_argument('--start-epoch', default=0, type=int, metavar='N',
                        help='manual epoch number (useful on restarts)')
    parser.add_argument('--batch-size', type=int, default=256,
                        help='mini-batch size (default: 256); the total over all GPUs of the node')
    parser.add_argument('--lr', '--learning-rate', default=1e-3, type=float, metavar='LR',
                        help='initial learning rate', dest='lr')
    parser.add_argument('--momentum', default=0.9, type=float, metavar='M',
                        help='momentum (used by --optimizer sgd)')
    parser.add_argument('--wd', '--weight-decay', default=1e-4, type=float, metavar='W',
                        help='weight decay (default: 1e-4; used by --optimizer sgd)', dest='weight_decay')
    parser.add_argument('--resume', default='', type=str, metavar='PATH',
                        help='path to latest checkpoint (default: none); "auto" = newest checkpoint_N.pth.tar in the '
                             'checkpoint directory')
    parser.add_argument('-e', '--evaluate', dest='evaluate', action='store_true',
                        help='evaluate model on validation set')
    parser.add_argument('--backend', type=str, default='nccl', help='Name of the backend to use.')
    parser.add_argument('--local-rank', '--local_rank', dest='local_rank', type=int, default=None,
                        help='set by torch.distributed.launch; also read from $LOCAL_RANK')
    parser.add_argument('-i', '--init-method', type=str, default='tcp://127.0.0.1:23456',
                        help='URL specifying how to initialize the package.')
    parser.add_argument('-s', '--world-size', type=int, default=1,
                        help='Number of processes participating in the job.')
    parser.add_argument('-r', '--rank', type=int, default=0, help='Rank of the current process.')
    parser.add_argument('--seed', default=None, type=int, help='seed for initializing training.')
    # ---- extensions ----
    parser.add_argument('--launcher', choices=('auto', 'spawn', 'launch'), default='auto',
                        help='spawn = torch.multiprocessing.spawn from this process; launch = one process per '
                             'rank started by torch.distributed.launch / torchrun; auto = detect')
    parser.add_argument('--optimizer', choices=('adam', 'sgd'), default='adam',
                        help='adam is what the reference runs (ref:191); sgd is its commented-out alternative')
    parser.add_argument('--arch', choices=('linear', 'cnn'), default='linear',
                        help='linear = the reference Net; cnn = the BASELINE.json MNIST CNN extension')
    parser.add_argument('--width', type=int, default=1, help='channel/width multiplier of --arch cnn')
    parser.add_argument('--precision', choices=('tf32', 'fp32x3'), default='tf32',
                        help='GEMM-shaped CNN layers: one tf32 tensor-core pass (default) or 3xTF32, which is '
                             'fp32-faithful at three times the tensor-core work (SURVEY 7.6); the reference Net is '
                             'fp32-faithful either way')
    parser.add_argument('--synthetic', action='store_true',
                        help='force the fixed-seed synthetic MNIST-shaped dataset (default: use idx files '
                             'under --root when present, synthetic otherwise)')
    parser.add_argument('--data-seed', type=int, default=1234, help='seed of the synthetic dataset')
    parser.add_argument('--train-size', type=int, default=None, help='truncate the training set (smoke tests)')
    parser.add_argument('--test-size', type=int, default=None, help='truncate the test set (smoke tests)')
    parser.add_argument('--device', choices=('auto', 'cuda', 'cpu'), default='auto')
    parser.add_argument('--comm', choices=('auto', 'peer', 'group'), default='auto',
                        help='gradient path: peer = our NVLink peer-memory kernels, group = torch.distributed')
    parser.add_argument('--store', choices=('torch', 'native'), default='torch',
                        help='key-value store for the peer-memory bootstrap: the process group\'s TCPStore, or '
                             'our native TCP store on the port after --init-method\'s')
    parser.add_argument('--fused', choices=('auto', 'on', 'off'), default='auto',
                        help='fused single-kernel training step for the reference Net')
    parser.add_argument('--data-placement', choices=('auto', 'staged', 'device', 'zerocopy'), default='auto',
                        help='fused Net path: staged = pinned ring + one H2D copy per step; device = training set '
                             'resident in HBM, gathered in-kernel; zerocopy = kernel gathers from pinned host memory; '
                             'auto = staged when the host has >= 12 cores per rank, else zerocopy')
    parser.add_argument('--checkpoint-dir', type=str, default=None, help='default: ./checkpoints (ref:264)')
    parser.add_argument('--comm-timeout', type=float, default=30.0,
                        help='seconds a peer-memory collective waits for a rank that does not show up before it '
                             'aborts with the missing rank\'s number (the reference hangs until the NCCL watchdog)')
    parser.add_argument('--print-topology', action='store_true',
                        help='rank 0 prints what the bootstrap found out about the GPU interconnect '
                             '(peer access matrix, NVLink/NVSwitch links, multicast support)')
    return parser


def detect_launcher(args, argv=None):
    """'launch' when started by torch.distributed.launch/torchrun, else 'spawn' (SURVEY D3)."""
    if args.launcher != 'auto':
        return args.launcher
    if 'LOCAL_RANK' in os.environ or 'RANK' in os.environ or 'TORCHELASTIC_RUN_ID' in os.environ:
        return 'launch'
    argv = sys.argv[1:] if argv is None else argv
    if any(a.split('=')[0] in ('--local-rank', '--local_rank') for a in argv):
        return 'launch'
    return 'spawn'


def resolve_local_rank(args):
    if args.local_rank is not None:
        return args.local_rank
    return int(os.environ.get('LOCAL_RANK', os.environ.get('RANK', 0)))


def seed_everything(seed):
    """ref:339-348, callable from every rank (in spawn mode the reference seeds only the parent)."""
    random.seed(seed)
    torch.manual_seed(seed)
    cudnn.deterministic = True
    torch.backends.cudnn.benchmark = False
    warnings.warn('You have chosen to seed training. '
                  'This will turn on the CUDNN deterministic setting, '
                  'which can slow down your training considerably! '
                  'You may see unexpected behavior when restarting '
                  'from checkpoints.')


def main(argv=None):
    from . import launch
    parser = build_parser()
    args = parser.parse_args(argv)
    print(args)

    if args.seed is not None:
        seed_everything(args.seed)

    mode = detect_launcher(args, argv)
    if torch.cuda.is_available() and args.device != 'cpu':
        ngpus = torch.cuda.device_count()
        if args.world_size > ngpus:
            raise SystemExit('--world-size {} exceeds the {} visible GPU(s); the reference requires '
                             'world_size == GPU count (ref:350-351)'.format(args.world_size, ngpus))
    if mode == 'launch':
        # way 1 (README:19): python -m torch.distributed.launch --nproc_per_node=N train.py --world-size N
        args.local_rank = resolve_local_rank(args)
        return launch.run_dist_launch(args)
    # way 2 (README:33): python train.py --world-size N
    return launch.demo_spawn(args.world_size, args)
