"""The two launch modes of the reference, selectable at run time.

* ``run_spawn`` / ``demo_spawn``  -- ref:273-276, ref:284-285: the parent calls
  ``mp.spawn`` and each child takes ``rank = proc_id``.
* ``run_dist_launch``              -- ref:278-281: an external launcher started
  one process per rank and ``rank = local_rank``.  As in the reference the
  launcher's ``MASTER_ADDR/MASTER_PORT`` are ignored; rendezvous uses
  ``--init-method``.
"""
import os

import torch.multiprocessing as mp


def run_spawn(proc_id, args):
    # single node, multiple GPUs: with torch.multiprocessing.spawn, rank = process id
    args.rank = proc_id
    if args.seed is not None:
        from .cli import seed_everything
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            seed_everything(args.seed)      # the reference leaves spawned children unseeded
    from .runner import run
    return run(args)


def run_dist_launch(args):
    # single node, multiple GPUs: with torch.distributed.launch, rank = local_rank
    args.rank = args.local_rank
    # torchrun-era launchers export TORCHELASTIC_USE_AGENT_STORE=True, which makes torch's
    # tcp:// rendezvous connect to the *launcher's* store as a client instead of letting
    # rank 0 host one at --init-method's port: every rank would wait forever on a port
    # nobody listens on.  The reference contract is "rendezvous at --init-method"
    # (ref:167-168), so opt out of the agent store for this process.
    os.environ.pop('TORCHELASTIC_USE_AGENT_STORE', None)
    from .runner import run
    return run(args)


def demo_spawn(nprocs, args):
    mp.spawn(run_spawn, args=(args,), nprocs=nprocs)
