"""Learning-rate schedule.

Parity: ``adjust_learning_rate`` ref:257-261 -- step decay by 10x every 10
epochs, recomputed from ``args.lr`` each epoch and written into every param
group (so a resumed checkpoint's stored ``lr`` is overwritten).
"""


def lr_for_epoch(base_lr: float, epoch: int) -> float:
    return base_lr * (0.1 ** (epoch // 10))


def adjust_learning_rate(optimizer, epoch, args):
    """Sets the learning rate to the initial LR decayed by 10 every 10 epochs."""
    lr = lr_for_epoch(args.lr, epoch)
    for group in optimizer.param_groups:
        group['lr'] = lr
    return lr
