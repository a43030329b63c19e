"""Host-side bootstrap of the (experimental) NVLS region: the multicast object travels as a file
descriptor, which needs SCM_RIGHTS over a Unix socket -- testable without a GPU."""
import datetime
import os

import torch.distributed as dist
import torch.multiprocessing as mp


def _fd_worker(rank, world, port, out_dir):
    from pytorch_distributed_mnist_b200.parallel import nvls
    store = dist.TCPStore('127.0.0.1', port, world, rank == 0, timeout=datetime.timedelta(seconds=60))
    if rank == 0:
        r, w = os.pipe()
        os.write(w, b'multicast-object')
        os.close(w)
        assert nvls.share_fd(store, 'pdm/test', 0, world, r) == r
        os.close(r)
    else:
        fd = nvls.share_fd(store, 'pdm/test', rank, world)
        # a duplicate of rank 0's descriptor: same pipe, so exactly one receiver gets the bytes
        os.set_blocking(fd, False)
        try:
            data = os.read(fd, 64)
        except BlockingIOError:
            data = b''
        os.close(fd)
        with open(os.path.join(out_dir, 'got{}'.format(rank)), 'wb') as f:
            f.write(data)
    nvls._wait_count(store, 'pdm/test/done', world, 30.0)
    # rank 0 hosts the store: it must outlive the other ranks' last poll
    if rank == 0:
        import time
        t0 = time.time()
        while int(store.add('pdm/test/left', 0)) < world - 1 and time.time() - t0 < 30:
            time.sleep(0.01)
    else:
        store.add('pdm/test/left', 1)


def test_descriptor_reaches_every_rank(tmp_path, free_port):
    world = 3
    mp.spawn(_fd_worker, args=(world, free_port, str(tmp_path)), nprocs=world)
    got = [open(str(tmp_path / 'got{}'.format(r)), 'rb').read() for r in range(1, world)]
    assert sorted(got) == [b'', b'multicast-object']        # both held the pipe; one drained it


def test_supported_never_raises_without_a_gpu():
    from pytorch_distributed_mnist_b200.parallel import nvls
    assert nvls.NvlsRegion.supported(0) in (False, True)     # False on this GPU-less box
