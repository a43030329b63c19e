"""MNIST storage: idx codec, on-disk discovery, and the synthetic stand-in.

The reference gets its data from ``torchvision.datasets.MNIST(download=True)``
and decodes every sample through PIL -> ``ToTensor`` -> ``Normalize`` on each
access (ref:132-138).  Here the dataset is two flat ``uint8``
arrays kept whole in memory (47 MB train): images ``[N, 784]`` and labels
``[N]``.  Normalisation ``(x/255 - 0.1307)/0.3081`` (ref:134) is applied per
*batch* -- on the host for the generic path, inside the CUDA kernels for the
fused path (the raw bytes are what crosses PCIe, 4x less than fp32).

There is no network on the target boxes, so when no idx files are found under
``root`` a fixed-seed synthetic set of the same shape is generated
(60,000 train / 10,000 test, ten learnable classes).  The same generator can
write idx files in torchvision's layout so the *unmodified* reference can be
benchmarked offline against identical bytes (``write_idx_tree``).
"""
from __future__ import annotations

import gzip
import os
import struct

import numpy as np

MEAN = 0.1307
STD = 0.3081
IMAGE_BYTES = 28 * 28
TRAIN_SIZE = 60000
TEST_SIZE = 10000

_FILES = {
    True: ('train-images-idx3-ubyte', 'train-labels-idx1-ubyte'),
    False: ('t10k-images-idx3-ubyte', 't10k-labels-idx1-ubyte'),
}


# ----------------------------------------------------------------------------
# idx codec
# ----------------------------------------------------------------------------
def _open(path):
    return gzip.open(path, 'rb') if path.endswith('.gz') else open(path, 'rb')


def read_idx(path):
    """Decode one idx file (ubyte payload) into a numpy array."""
    with _open(path) as f:
        raw = f.read()
    zero, dtype_code, ndim = struct.unpack('>HBB', raw[:4])
    if zero != 0 or dtype_code != 0x08:
        raise ValueError('{}: not a ubyte idx file (magic {!r})'.format(path, raw[:4]))
    dims = struct.unpack('>' + 'I' * ndim, raw[4:4 + 4 * ndim])
    data = np.frombuffer(raw, dtype=np.uint8, offset=4 + 4 * ndim)
    expect = int(np.prod(dims))
    if data.size != expect:
        raise ValueError('{}: payload {} bytes, header says {}'.format(path, data.size, expect))
    return data.reshape(dims).copy()


def write_idx(path, array):
    array = np.ascontiguousarray(array, dtype=np.uint8)
    header = struct.pack('>HBB', 0, 0x08, array.ndim) + struct.pack('>' + 'I' * array.ndim, *array.shape)
    tmp = path + '.tmp.{}'.format(os.getpid())
    with open(tmp, 'wb') as f:
        f.write(header)
        f.write(array.tobytes())
    os.replace(tmp, path)


def raw_folder(root):
    """torchvision's layout: ``<root>/MNIST/raw``."""
    return os.path.join(root, 'MNIST', 'raw')


def find_idx(root, train):
    """Return (images_path, labels_path) if both exist under ``root`` in either
    torchvision's layout or directly in ``root``; else None."""
    img, lab = _FILES[train]
    for folder in (raw_folder(root), root):
        for ext in ('', '.gz'):
            pi, pl = os.path.join(folder, img + ext), os.path.join(folder, lab + ext)
            if os.path.isfile(pi) and os.path.isfile(pl):
                return pi, pl
    return None


# ----------------------------------------------------------------------------
# synthetic data
# ----------------------------------------------------------------------------
def _prototypes(rng):
    """Ten smooth 28x28 class templates in [0, 1]."""
    yy, xx = np.mgrid[0:28, 0:28].astype(np.float32)
    protos = np.zeros((10, 28, 28), dtype=np.float32)
    for c in range(10):
        for _ in range(4):
            cy, cx = rng.uniform(5, 23, size=2)
            sy, sx = rng.uniform(1.5, 4.5, size=2)
            protos[c] += np.exp(-(((yy - cy) / sy) ** 2 + ((xx - cx) / sx) ** 2))
        protos[c] /= protos[c].max()
    return protos


def synthetic_mnist(train=True, seed=1234, size=None):
    """Fixed-seed MNIST-shaped set: (uint8 images [N, 28, 28], uint8 labels [N]).

    Each class is a smooth template blended 50/50 with a background common to
    all classes, plus heavy pixel noise: a linear model tops out near 94%
    (MNIST-like), so accuracy and loss curves are informative for the
    equivalence tests.
    """
    n = size if size is not None else (TRAIN_SIZE if train else TEST_SIZE)
    proto_rng = np.random.RandomState(seed)          # same templates for train and test
    protos = _prototypes(proto_rng)
    common = _prototypes(np.random.RandomState(seed + 77)).mean(axis=0)
    protos = 0.5 * protos + 0.5 * common[None]
    rng = np.random.RandomState(seed + (1 if train else 2))
    labels = rng.randint(0, 10, size=n).astype(np.uint8)
    images = np.empty((n, 28, 28), dtype=np.uint8)
    chunk = 10000
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        amp = rng.uniform(0.4, 1.0, size=(e - s, 1, 1)).astype(np.float32)
        noise = rng.normal(0.0, 0.35, size=(e - s, 28, 28)).astype(np.float32)
        x = protos[labels[s:e]] * amp + noise
        images[s:e] = np.clip(x * 255.0, 0, 255).astype(np.uint8)
    return images, labels


def write_idx_tree(root, seed=1234, train_size=None, test_size=None):
    """Materialise the synthetic set as idx files where torchvision looks for
    them, so ``datasets.MNIST(root, download=True)`` finds them and skips the
    download.  Idempotent; safe under concurrent callers (atomic renames)."""
    folder = raw_folder(root)
    os.makedirs(folder, exist_ok=True)
    for train, size in ((True, train_size), (False, test_size)):
        img, lab = _FILES[train]
        pi, pl = os.path.join(folder, img), os.path.join(folder, lab)
        if os.path.isfile(pi) and os.path.isfile(pl):
            continue
        images, labels = synthetic_mnist(train, seed=seed, size=size)
        write_idx(pi, images)
        write_idx(pl, labels)
    return folder


def load_mnist(root, train=True, synthetic=None, seed=1234, size=None):
    """Return (images uint8 [N, 784], labels uint8 [N], source string).

    ``synthetic=None`` means auto: real idx files if present under ``root``,
    otherwise the synthetic set.  ``synthetic=True`` forces synthetic.
    """
    found = None if synthetic else find_idx(root, train)
    if found is not None:
        images = read_idx(found[0])
        labels = read_idx(found[1])
        source = 'idx:' + found[0]
    elif synthetic is False:
        raise FileNotFoundError(
            'no MNIST idx files under {!r} and downloads are unavailable offline; '
            'pass --synthetic or place the four idx files in {!r}'.format(root, raw_folder(root)))
    else:
        images, labels = synthetic_mnist(train, seed=seed, size=size)
        source = 'synthetic(seed={})'.format(seed)
    if size is not None:
        images, labels = images[:size], labels[:size]
    images = np.ascontiguousarray(images.reshape(images.shape[0], IMAGE_BYTES))
    return images, np.ascontiguousarray(labels.astype(np.uint8)), source
