import sys, time, os
sys.path.insert(0, '/root/repo')
import torch
import pytorch_distributed_mnist_b200 as pdm
from pytorch_distributed_mnist_b200.ops.engine import NetEngine
dev = torch.device('cuda', 0)
net = pdm.Net().to(dev); opt = pdm.FusedAdam(net.parameters(), lr=1e-3)
B = 512
eng = NetEngine(opt.flat, opt, max_batch=B)
nb = 512
x = torch.randint(0, 256, (nb * B, 784), dtype=torch.uint8, device=dev)
y = torch.randint(0, 10, (nb * B,), dtype=torch.uint8, device=dev)
eng.train_device(x, y, B, 5)
for label, k, sleep in (('back-to-back', 20, 0), ('back-to-back', 20, 0), ('back-to-back', 1, 0), ('back-to-back', 2, 0), ('after 50 ms idle', 20, 0.05), ('after 50 ms idle', 20, 0.05),
                        ('back-to-back', 200, 0), ('after 50 ms idle', 200, 0.05), ('back-to-back', 2000, 0), ('back-to-back', 20, 0)):
    if sleep: time.sleep(sleep)
    ms = eng.train_device(x, y, B, k)
    print('{:18s} K={:5d}: {:8.2f} us total, {:6.2f} us/step'.format(label, k, ms * 1e3, ms * 1e3 / k), flush=True)
