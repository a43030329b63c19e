"""pytorch_distributed_mnist_b200 -- a Blackwell-native data-parallel MNIST trainer
with the capabilities and command line of flybirdtian/pytorch_distributed_mnist.

Public surface (names match the reference module where one exists)::

    Net, CNN                      models
    MNISTDataLoader               data
    Trainer, Average, Accuracy    train / evaluate loop and metrics
    DistributedDataParallel       gradient path (peer-memory kernels or torch.distributed)
    FusedAdam, FusedSGD           flat fused optimizers, torch.optim-compatible state
    run, main                     per-process job and CLI
    save_checkpoint, adjust_learning_rate, distributed_is_initialized
"""
from .metrics import Average, Accuracy
from .schedule import adjust_learning_rate
from .checkpoint import save_checkpoint, load_checkpoint
from .data.loader import MNISTDataLoader, distributed_is_initialized
from .models import Net, CNN, build_model
from .optim import FusedAdam, FusedSGD, build_optimizer
from .parallel.ddp import DistributedDataParallel
from .trainer import Trainer
from .runner import run
from .launch import run_spawn, run_dist_launch, demo_spawn
from .cli import main, build_parser

__version__ = '0.1.0'

__all__ = ['Average', 'Accuracy', 'adjust_learning_rate', 'save_checkpoint', 'load_checkpoint',
           'MNISTDataLoader', 'distributed_is_initialized', 'Net', 'CNN', 'build_model',
           'FusedAdam', 'FusedSGD', 'build_optimizer', 'DistributedDataParallel', 'Trainer',
           'run', 'run_spawn', 'run_dist_launch', 'demo_spawn', 'main', 'build_parser']
