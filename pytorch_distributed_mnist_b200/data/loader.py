"""``MNISTDataLoader`` -- same constructor and iteration contract as the
reference's (ref:129-161), rebuilt around a flat in-memory uint8 dataset.

What is kept:

* ctor ``(root, batch_size, num_workers, train=True)``; ``set_sample_epoch``.
* a distributed sampler only when ``train`` and the process group is up
  (ref:142-144); shuffle only when training without a sampler (ref:146-149);
  the test set is never sharded: every rank evaluates all 10,000 images in
  file order.
* iteration yields ``(float32 [B,1,28,28] normalised, int64 [B])`` CPU
  tensors, last batch ragged (``drop_last`` unset).
* ``ShardSampler`` reproduces ``torch.utils.data.DistributedSampler`` index
  for index (seed 0 + epoch, pad by repeating the head of the permutation,
  rank-strided subsample), so a run here sees the same samples in the same
  order as stock DDP.  That is what makes the end-to-end equivalence test
   exact.

What is new: no per-sample PIL decode and no worker *processes*.  A batch is a
row gather from one uint8 array.  ``raw_batches()`` hands the fused CUDA path
the raw bytes + labels (normalisation happens in-kernel); the native engine
(``ops.engine``) takes the whole index plan for an epoch and runs the gather in
C++ worker threads into a pinned ring (the N14 equivalent).
"""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.distributed as dist

from . import mnist


def distributed_is_initialized():
    """ref:21-25."""
    if dist.is_available():
        if dist.is_initialized():
            return True
    return False


class ShardSampler(object):
    """Index-exact restatement of ``DistributedSampler(dataset)`` defaults
    (shuffle=True, seed=0, drop_last=False)."""

    def __init__(self, length, num_replicas=None, rank=None, shuffle=True, seed=0):
        if num_replicas is None:
            num_replicas = dist.get_world_size() if distributed_is_initialized() else 1
        if rank is None:
            rank = dist.get_rank() if distributed_is_initialized() else 0
        if not 0 <= rank < num_replicas:
            raise ValueError('rank {} outside [0, {})'.format(rank, num_replicas))
        self.length = length
        self.num_replicas = num_replicas
        self.rank = rank
        self.shuffle = shuffle
        self.seed = seed
        self.epoch = 0
        self.num_samples = math.ceil(length / num_replicas)
        self.total_size = self.num_samples * num_replicas

    def set_epoch(self, epoch):
        self.epoch = epoch

    def indices(self, epoch=None):
        """This rank's sample order for ``epoch`` (default: the epoch set by ``set_epoch``).  Pure:
        safe to call for several epochs from several threads."""
        epoch = self.epoch if epoch is None else epoch
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + epoch)
            idx = torch.randperm(self.length, generator=g)
        else:
            idx = torch.arange(self.length)
        pad = self.total_size - idx.numel()
        if pad > 0:
            reps = math.ceil(pad / idx.numel())
            idx = torch.cat([idx, idx.repeat(reps)[:pad]])
        return idx[self.rank:self.total_size:self.num_replicas].contiguous()

    def __iter__(self):
        return iter(self.indices().tolist())

    def __len__(self):
        return self.num_samples


class MNISTDataLoader(object):

    def __init__(self, root, batch_size, num_workers=0, train=True,
                 synthetic=None, seed=1234, size=None, samples_per_epoch=None):
        if batch_size < 1:
            raise ValueError('per-process batch size must be >= 1 (got {}); '
                             '--batch-size is the node total and is divided by the GPU count'
                             .format(batch_size))
        images, labels, source = mnist.load_mnist(root, train=train, synthetic=synthetic,
                                                  seed=seed, size=size)
        self.images = torch.from_numpy(images)        # uint8 [N, 784]
        self.labels = torch.from_numpy(labels)        # uint8 [N]
        self.source = source
        self.train = train
        self.batch_size = batch_size
        self.num_workers = num_workers
        # Optional fixed pass length (per process): a pass then walks as many successive epoch
        # permutations as needed.  Fixed-length "virtual epochs" for schedules and benchmarks.
        self.samples_per_epoch = samples_per_epoch
        self._plan = None

        self.sampler = None
        if train and distributed_is_initialized():
            self.sampler = ShardSampler(len(self.labels))
        if train:
            self.shuffle = (self.sampler is None)
        else:
            self.shuffle = False

    # -- reference surface --------------------------------------------------
    def set_sample_epoch(self, epoch=0):
        if self.train and self.sampler is not None:
            self.sampler.set_epoch(epoch)
            # The epoch's index plan depends only on (seed, epoch): build it now, once, instead of
            # inside the training pass (the permutation of 60,000 indices costs more than twenty
            # fused steps).  ``epoch_indices()`` hands this plan out.
            self._plan = (epoch, self._build_plan())

    @property
    def dataset(self):
        return self

    def num_samples(self):
        if self.samples_per_epoch is not None:
            return int(self.samples_per_epoch)
        return len(self.sampler) if self.sampler is not None else len(self.labels)

    def __len__(self):
        return math.ceil(self.num_samples() / self.batch_size)

    # -- index plan ---------------------------------------------------------
    def _one_epoch(self, epoch_offset=0):
        if self.sampler is not None:
            return self.sampler.indices(self.sampler.epoch + epoch_offset)
        if self.shuffle:
            return torch.randperm(len(self.labels))
        return torch.arange(len(self.labels))

    def epoch_indices(self):
        """The sample order for one pass, as an int64 tensor."""
        plan = self._plan
        if plan is not None and self.sampler is not None and plan[0] == self.sampler.epoch:
            return plan[1]
        return self._build_plan()

    def _build_plan(self):
        if self.samples_per_epoch is None:
            return self._one_epoch()
        want = int(self.samples_per_epoch)
        per_part = self.sampler.num_samples if self.sampler is not None else len(self.labels)
        n_parts = max(1, -(-want // max(1, per_part)))
        # (building the permutations on a thread pool was tried: slower, torch.randperm is already the whole cost)
        parts = [self._one_epoch(k) for k in range(n_parts)]
        return torch.cat(parts)[:want].contiguous()

    # -- iteration ----------------------------------------------------------
    def raw_batches(self):
        """Yield ``(uint8 [B, 784], uint8 [B])`` -- bytes as stored."""
        idx = self.epoch_indices()
        for s in range(0, idx.numel(), self.batch_size):
            sel = idx[s:s + self.batch_size]
            yield self.images.index_select(0, sel), self.labels.index_select(0, sel)

    @staticmethod
    def normalize(images_u8):
        """ToTensor + Normalize((0.1307,), (0.3081,)) for a whole batch."""
        x = images_u8.to(torch.float32).div_(255.0).sub_(mnist.MEAN).div_(mnist.STD)
        return x.view(-1, 1, 28, 28)

    def __iter__(self):
        for img, lab in self.raw_batches():
            yield self.normalize(img), lab.to(torch.int64)


def as_numpy_indices(idx):
    return np.ascontiguousarray(idx.numpy().astype(np.int64))
