#!/usr/bin/env python
"""Staged, verbose exercise of the native engine (run with PDM_ENGINE_DEBUG=1 under `timeout`)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def say(*a):
    print('[debug_engine %.2f]' % time.time(), *a, flush=True)


import torch  # noqa: E402

say('torch imported')
import pytorch_distributed_mnist_b200 as pdm  # noqa: E402
from pytorch_distributed_mnist_b200.data import mnist  # noqa: E402
from pytorch_distributed_mnist_b200.ops.engine import NetEngine  # noqa: E402

torch.manual_seed(0)
net = pdm.Net().cuda()
opt = pdm.FusedAdam(net.parameters(), lr=1e-3)
say('model + optimizer built')
images, labels = mnist.synthetic_mnist(True, size=2048)
images = torch.from_numpy(images.reshape(-1, 784))
labels = torch.from_numpy(labels)
eng = NetEngine(opt.flat, opt, max_batch=128, slots=4, workers=2)
say('engine created')
eng.set_dataset(images, labels)
say('dataset set')
ms = eng.train_device(images.cuda(), labels.cuda(), 128, 6)
say('device path: 6 steps in %.3f ms' % ms)
steps, ms = eng.train(torch.arange(640), 128)
say('host path: %d steps in %.3f ms, losses %s' % (steps.shape[0], ms, steps[:, 0].round(3).tolist()))
steps, ms = eng.train(torch.randperm(2048), 100)
say('host path shuffled: %d steps in %.3f ms (%.1f us/step)' % (steps.shape[0], ms, ms * 1e3 / steps.shape[0]))
ms = eng.evaluate(torch.arange(2048), 128)
say('eval: %.3f ms, metrics %s' % (ms, eng.metrics()))
eng.close()
say('closed')
