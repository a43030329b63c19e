"""The C ABI validates its arguments before it touches the device, so the error paths can be
exercised on a box without a GPU: every call below must come back with a non-zero status and a
message in pdm_last_error(), never crash."""
import ctypes

import pytest

from pytorch_distributed_mnist_b200.ops import _lib

c_void_p = ctypes.c_void_p


@pytest.fixture(scope='module')
def lib():
    return _lib.lib()


def _err(lib):
    return lib.pdm_last_error().decode('utf-8', 'replace')


def test_cnn_executor_rejects_bad_shapes(lib):
    from pytorch_distributed_mnist_b200.ops import cnn_engine  # noqa: F401  (registers argtypes)
    h = c_void_p()
    assert lib.pdm_cnn_create(ctypes.byref(h), 0, 9, 64) != 0 and 'width' in _err(lib)
    assert lib.pdm_cnn_create(ctypes.byref(h), 0, 1, 0) != 0
    assert lib.pdm_cnn_create(None, 0, 1, 64) != 0
    assert lib.pdm_cnn_destroy(None) == 0                       # destroying nothing is fine


def test_nvls_region_rejects_bad_teams(lib):
    from pytorch_distributed_mnist_b200.parallel import nvls  # noqa: F401
    h = c_void_p()
    assert lib.pdm_nvls_create(ctypes.byref(h), 0, 1, 0, 4096) != 0 and 'world' in _err(lib)      # a team of one
    assert lib.pdm_nvls_create(ctypes.byref(h), 3, 2, 0, 4096) != 0                                # rank outside the team
    assert lib.pdm_nvls_create(None, 0, 2, 0, 4096) != 0
    assert lib.pdm_nvls_destroy(None) == 0
    assert lib.pdm_nvls_supported(0) in (0, 1, -1)


def test_communicator_rejects_bad_ranks(lib):
    h = c_void_p()
    assert lib.pdm_comm_create(ctypes.byref(h), 0, 9, 0, 4096) != 0 and 'world' in _err(lib)       # more than PDM_MAX_RANKS
    assert lib.pdm_comm_create(ctypes.byref(h), 2, 2, 0, 4096) != 0 and 'rank' in _err(lib)
    assert lib.pdm_comm_create(None, 0, 1, 0, 4096) != 0


def test_fused_step_rejects_null_arguments(lib):
    assert lib.pdm_net_train_step(None, None) != 0 and 'NULL' in _err(lib)


def test_bucket_assignment_matches_python_twin(lib):
    from pytorch_distributed_mnist_b200.parallel.ddp import assign_buckets
    sizes = [288, 32, 18432, 64, 1179648, 128, 1280, 10]        # the CNN, width 1
    got = assign_buckets(sizes)
    assert [sorted(b) for b in got] == [[4, 5, 6, 7], [0, 1, 2, 3]]
    assert sum(len(b) for b in assign_buckets(list(range(1, 200)), cap_bytes=1024)) == 199


def test_ctypes_structs_match_the_header(lib):
    """Every struct that crosses the C ABI has the size this build of the library compiled it with
    (a field added to include/pdmnist.h but not to the ctypes mirror would silently shift everything after it)."""
    from pytorch_distributed_mnist_b200.ops import cnn_engine, gemm
    from pytorch_distributed_mnist_b200.parallel import reducer
    lib.pdm_abi_sizeof.restype = ctypes.c_size_t
    lib.pdm_abi_sizeof.argtypes = [ctypes.c_int]
    mirrors = [_lib.Hyper, _lib.Peers, _lib.NetStepArgs, _lib.NetTcArgs, gemm.GemmArgs, cnn_engine.CnnParams,
               cnn_engine.CnnPlan, reducer.ReducerConfig]
    for which, cls in enumerate(mirrors):
        assert lib.pdm_abi_sizeof(which) == ctypes.sizeof(cls), cls.__name__
    assert lib.pdm_abi_sizeof(len(mirrors)) == 0
