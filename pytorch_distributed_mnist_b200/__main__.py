"""``python -m pytorch_distributed_mnist_b200 ...`` == the reference's ``python multi_proc_single_gpu.py ...``
(ref:288-359; README:11-33): same flags, both launch modes."""
from .cli import main

if __name__ == '__main__':
    main()
