"""``CNN`` -- the BASELINE.json "MNIST CNN" extension.

The reference defines no CNN; this is the canonical PyTorch-examples MNIST
net, chosen and recorded as an assumption in SURVEY 7.4::

    conv1 1->32 3x3, ReLU, conv2 32->64 3x3, ReLU, max-pool 2, dropout 0.25,
    fc1 9216->128, ReLU, dropout 0.5, fc2 128->10            (1,199,882 params)

``width`` multiplies the channel counts and the fc1 width (``--width 8`` is
BASELINE config 4: 256/512 channels, fc1 73,728->1,024, ~78 M parameters,
~310 MB of gradients -- the large-bucket / NVLS stress case).

On CUDA every layer dispatches to the hand-written sm_100a kernels in
``ops`` (tcgen05 tf32 implicit-GEMM conv, fused bias+ReLU(+max-pool), Philox
dropout with the mask recomputed in backward, tcgen05 linear).  On CPU the
same module runs stock PyTorch ops, which is also the numerics oracle.
"""
import torch.nn as nn
import torch.nn.functional as F


class CNN(nn.Module):

    def __init__(self, width=1, p1=0.25, p2=0.5):
        super(CNN, self).__init__()
        self.width = int(width)
        c1, c2, h = 32 * self.width, 64 * self.width, 128 * self.width
        self.conv1 = nn.Conv2d(1, c1, 3, 1)
        self.conv2 = nn.Conv2d(c1, c2, 3, 1)
        self.fc1 = nn.Linear(c2 * 12 * 12, h)
        self.fc2 = nn.Linear(h, 10)
        self.p1, self.p2 = p1, p2

    def forward(self, x):
        x = x.view(x.size(0), 1, 28, 28)
        if x.is_cuda:
            from ..ops import cnn_ops
            if cnn_ops.available():
                return cnn_ops.cnn_forward(self, x)
        x = F.relu(self.conv1(x))
        x = F.relu(self.conv2(x))
        x = F.max_pool2d(x, 2)
        x = F.dropout(x, self.p1, self.training)
        x = x.flatten(1)
        x = F.relu(self.fc1(x))
        x = F.dropout(x, self.p2, self.training)
        return self.fc2(x)
