"""Running metrics with the reference's public surface and print formats.

Parity targets (``ref:N`` = /root/reference/multi_proc_single_gpu.py line N):

* ``Average``  -- ref:28-43.  Sample-weighted running mean, ``'{:.6f}'`` format.
* ``Accuracy`` -- ref:46-65.  Running top-1, ``'{:.2f}%'`` format.

What is new here: both classes can also be fed *device-side accumulators*
(``update_from_device``) so the hot loop never calls ``.item()``.  The
reference pays two full stream syncs per step for its metrics (ref:94-95); the
B200 trainer keeps sums on the device and reads them back once per epoch.
"""
from __future__ import annotations

import torch


class Average(object):
    """Sample-weighted mean of a scalar (loss)."""

    def __init__(self):
        self.sum = 0
        self.count = 0

    def __str__(self):
        return '{:.6f}'.format(self.average)

    @property
    def average(self):
        # Same failure mode as the reference: ZeroDivisionError when never updated.
        return self.sum / self.count

    def update(self, value, number):
        """``value`` is the batch-mean, ``number`` the batch size."""
        self.sum += value * number
        self.count += number

    def update_sum(self, total, number):
        """Add an already sample-weighted sum (one epoch readback)."""
        self.sum += total
        self.count += number


class Accuracy(object):
    """Running top-1 accuracy."""

    def __init__(self):
        self.correct = 0
        self.count = 0

    def __str__(self):
        return '{:.2f}%'.format(self.accuracy * 100)

    @property
    def accuracy(self):
        return self.correct / self.count

    @torch.no_grad()
    def update(self, output, target):
        pred = output.argmax(dim=1)
        correct = pred.eq(target).sum().item()
        self.correct += correct
        self.count += output.size(0)

    def update_counts(self, correct, number):
        """Add a pre-counted number of hits (one epoch readback)."""
        self.correct += int(correct)
        self.count += int(number)


class DeviceMetrics(object):
    """Device-resident ``[loss_sum, correct, samples]`` accumulator.

    Kernels add into it (fp64 loss sum, int64 counters are emulated with two
    fp64 slots so one tiny buffer serves all three); ``drain`` performs the one
    device-to-host read of an epoch and folds it into Average/Accuracy.
    """

    def __init__(self, device):
        self.buf = torch.zeros(4, dtype=torch.float64, device=device)

    def zero_(self):
        self.buf.zero_()

    def add_torch(self, loss_mean, output, target):
        """Generic (non-fused) path: accumulate without a host sync."""
        n = output.size(0)
        self.buf[0] += loss_mean.detach().double() * n
        self.buf[1] += (output.detach().argmax(dim=1) == target).sum().double()
        self.buf[2] += n

    def drain(self, avg: Average, acc: Accuracy):
        loss_sum, correct, samples, _ = self.buf.tolist()  # the one sync
        samples = int(round(samples))
        if samples:
            avg.update_sum(loss_sum, samples)
            acc.update_counts(int(round(correct)), samples)
        self.zero_()
        return avg, acc
