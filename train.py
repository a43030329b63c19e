#!/usr/bin/env python
"""Entry script: drop-in for the reference's ``multi_proc_single_gpu.py``.

    # way 1 (README:19), no source edit needed:
    python -m torch.distributed.launch --nproc_per_node=4 train.py --world-size 4 --workers 4
    # way 2 (README:33):
    python train.py --world-size 4
    # evaluate / resume (README:50-60):
    python train.py --world-size 1 --evaluate --resume checkpoints/model_best.pth.tar
"""
from pytorch_distributed_mnist_b200.cli import main

if __name__ == '__main__':
    main()
