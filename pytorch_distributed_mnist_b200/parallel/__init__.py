from .flat import FlatState, flat_of
from .peer import PeerComm, GroupComm
from .ddp import DistributedDataParallel, assign_buckets

__all__ = ['FlatState', 'flat_of', 'PeerComm', 'GroupComm', 'DistributedDataParallel', 'assign_buckets']
