from .net import Net
from .cnn import CNN


def build_model(arch='linear', width=1):
    """``--arch linear`` is the reference ``Net``; ``--arch cnn`` the extension."""
    if arch == 'linear':
        return Net()
    if arch == 'cnn':
        return CNN(width=width)
    raise ValueError('unknown --arch {!r} (choose linear or cnn)'.format(arch))


__all__ = ['Net', 'CNN', 'build_model']
