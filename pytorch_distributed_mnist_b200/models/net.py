"""``Net`` -- the reference model: one ``Linear(784, 10)`` on the flattened
image (ref:119-126).  7,850 fp32 parameters; ``state_dict``
keys ``fc.weight`` (10, 784) and ``fc.bias`` (10,), so reference checkpoints
load unchanged.

On a CUDA device the forward runs the hand-written sm_100a linear kernel
(``ops.linear``); the training hot loop does not call ``forward`` at all -- the
trainer hands ``Net`` to the fused step kernel (forward + loss + backward +
all-reduce + optimizer in one launch, ``ops.fused_step``).
"""
import torch.nn as nn
import torch.nn.functional as F


class Net(nn.Module):

    IN_FEATURES = 784
    NUM_CLASSES = 10

    def __init__(self):
        super(Net, self).__init__()
        self.fc = nn.Linear(self.IN_FEATURES, self.NUM_CLASSES)

    def forward(self, x):
        x = x.view(x.size(0), -1)
        if x.is_cuda:
            from ..ops import linear as _linear
            return _linear.linear(x, self.fc.weight, self.fc.bias)
        return F.linear(x, self.fc.weight, self.fc.bias)
