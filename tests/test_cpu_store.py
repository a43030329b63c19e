"""Native TCP rendezvous store (csrc/store.cu): multi-process, CPU only."""
import multiprocessing as mp
import socket
import time

import pytest


def _port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    from pytorch_distributed_mnist_b200.parallel.store import NativeStore
    try:
        if rank != 0:
            time.sleep(0.3)                       # clients may come up before the server: connect retries
        st = NativeStore('127.0.0.1', port, is_server=(rank == 0), timeout_s=30)
        st.set('handle/{}'.format(rank), bytes([rank]) * 64)
        got = [st.get('handle/{}'.format(r)) for r in range(world)]          # blocks until published
        assert got == [bytes([r]) * 64 for r in range(world)]
        n = st.add('arrived', 1)
        assert 1 <= n <= world
        while st.add('arrived', 0) < world:
            time.sleep(0.01)
        st.set('empty', b'')
        assert st.get('empty') == b''
        if rank == 0:
            time.sleep(0.5)                       # keep the server alive until everyone is done
        st.close()
        q.put((rank, 'ok'))
    except Exception as e:  # noqa: BLE001
        q.put((rank, repr(e)))


@pytest.mark.parametrize('world', [1, 4])
def test_store_rendezvous(world):
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = dict(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(30)
    assert results == {r: 'ok' for r in range(world)}


def test_store_get_times_out():
    from pytorch_distributed_mnist_b200.ops import _lib
    from pytorch_distributed_mnist_b200.parallel.store import NativeStore
    st = NativeStore('127.0.0.1', _port(), is_server=True, timeout_s=1.0)
    t0 = time.time()
    with pytest.raises(_lib.NativeError, match='timed out'):
        st.get('never-published')
    assert 0.5 < time.time() - t0 < 10
    st.close()


def test_store_concurrent_clients():
    """Eight clients hammering one server (set / add / blocking get of a neighbour's key): the
    counter must be exact and every value intact."""
    import threading
    from pytorch_distributed_mnist_b200.parallel.store import NativeStore
    port = _port()
    server = NativeStore('127.0.0.1', port, is_server=True, timeout_s=30)
    errs, clients, rounds = [], 8, 120

    def client(i):
        try:
            st = NativeStore('127.0.0.1', port, is_server=False, timeout_s=30)
            for k in range(rounds):
                st.set('k/{}/{}'.format(i, k), bytes([i]) * (k % 97 + 1))
                st.add('ctr', 1)
                if k % 10 == 0:
                    j = (i + 1) % clients
                    assert st.get('k/{}/{}'.format(j, k)) == bytes([j]) * (k % 97 + 1)
            st.close()
        except Exception as e:  # noqa: BLE001
            errs.append(repr(e))

    threads = [threading.Thread(target=client, args=(i,)) for i in range(clients)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(90)
    assert not errs and not any(t.is_alive() for t in threads), errs[:3]
    assert server.add('ctr', 0) == clients * rounds
    server.close()
