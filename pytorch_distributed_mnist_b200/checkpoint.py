"""Checkpoint I/O in the reference's on-disk layout.

Parity: ``save_checkpoint`` ref:263-271 and the resume block ref:196-214.

Layout (unchanged, so reference checkpoints resume here and vice versa)::

    {'epoch': int, 'state_dict': {'module.<name>': tensor, ...},
     'best_acc': float, 'optimizer': torch.optim.Adam-style state_dict}

Files: ``checkpoints/checkpoint_{epoch}.pth.tar`` (torch zip format despite the
name) and ``checkpoints/model_best.pth.tar``.  What differs from the
reference: the write is atomic (tmp file + rename) so a rank that dies while
saving never leaves a truncated checkpoint behind, and the directory can be
overridden (the reference hard-codes a cwd-relative ``checkpoints``).
"""
from __future__ import annotations

import os
import shutil

import torch

CHECKPOINT_DIR = 'checkpoints'


def checkpoint_path(epoch, chk_dir=CHECKPOINT_DIR):
    return os.path.join(chk_dir, 'checkpoint_{}.pth.tar'.format(epoch))


def best_path(chk_dir=CHECKPOINT_DIR):
    return os.path.join(chk_dir, 'model_best.pth.tar')


def save_checkpoint(state, is_best, epoch, chk_dir=CHECKPOINT_DIR):
    os.makedirs(chk_dir, exist_ok=True)
    filename = checkpoint_path(epoch, chk_dir)
    tmp = filename + '.tmp.{}'.format(os.getpid())
    torch.save(state, tmp)
    os.replace(tmp, filename)
    if is_best:
        shutil.copyfile(filename, best_path(chk_dir))
    return filename


def latest_checkpoint(chk_dir=CHECKPOINT_DIR):
    """Path of ``checkpoint_{epoch}.pth.tar`` with the highest epoch in ``chk_dir``, or None.
    Used by ``--resume auto``: the reference's recovery story is "rerun with --resume PATH"
    (README:50-52); this saves looking the path up after a crash."""
    import re
    best, best_epoch = None, -1
    if os.path.isdir(chk_dir):
        for name in os.listdir(chk_dir):
            m = re.fullmatch(r'checkpoint_(\d+)\.pth\.tar', name)
            if m and int(m.group(1)) > best_epoch:
                best, best_epoch = os.path.join(chk_dir, name), int(m.group(1))
    return best


def load_checkpoint(path, map_location=None):
    """``torch.load`` with the safe (weights_only) default; reference
    checkpoints hold only tensors and primitives so this needs no override."""
    return torch.load(path, map_location=map_location)


def _strip_prefix(key, prefix='module.'):
    return key[len(prefix):] if key.startswith(prefix) else key


def adapt_state_dict_keys(state_dict, model):
    """Accept both ``module.``-prefixed (saved from a DDP wrapper, ref:252) and
    bare keys, whichever the target model wants."""
    want = set(model.state_dict().keys())
    if set(state_dict.keys()) == want:
        return state_dict
    wants_prefix = any(k.startswith('module.') for k in want)
    out = type(state_dict)()
    for k, v in state_dict.items():
        bare = _strip_prefix(k)
        out['module.' + bare if wants_prefix else bare] = v
    return out


def make_state(epoch, model, best_acc, optimizer):
    """The dict the reference builds inline at ref:250-255."""
    return {
        'epoch': epoch + 1,
        'state_dict': model.state_dict(),
        'best_acc': best_acc,
        'optimizer': optimizer.state_dict(),
    }
