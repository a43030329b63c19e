from .loader import MNISTDataLoader, ShardSampler, distributed_is_initialized
from . import mnist

__all__ = ['MNISTDataLoader', 'ShardSampler', 'distributed_is_initialized', 'mnist']
