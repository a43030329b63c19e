import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

REFERENCE_DIR = '/root/reference'


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box with `pytest -m gpu`)')
    config.addinivalue_line('markers', 'multigpu: needs at least two CUDA devices')


def pytest_collection_modifyitems(config, items):
    import torch
    ngpu = torch.cuda.device_count() if torch.cuda.is_available() else 0
    for item in items:
        if 'gpu' in item.keywords and ngpu == 0:
            item.add_marker(pytest.mark.skip(reason='no CUDA device'))
        if 'multigpu' in item.keywords and ngpu < 2:
            item.add_marker(pytest.mark.skip(reason='needs >= 2 CUDA devices'))


@pytest.fixture(autouse=True)
def fp32_oracles():
    """Every stock-PyTorch result a test compares against is computed in true fp32: cuDNN convolutions
    default to TF32 (matmul does not), which would make a tf32 kernel look exact against a tf32 oracle."""
    import torch
    prev = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = prev


@pytest.fixture(scope='session')
def reference_module():
    """The reference's own module, imported read-only as the oracle."""
    path = os.path.join(REFERENCE_DIR, 'multi_proc_single_gpu.py')
    alt = os.path.join(ROOT, 'baseline', '_ref', 'multi_proc_single_gpu.py')
    for p in (path, alt):
        if os.path.isfile(p):
            import importlib.util
            spec = importlib.util.spec_from_file_location('multi_proc_single_gpu_ref', p)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            return mod
    pytest.skip('reference module not available')


_port = [29700 + (os.getpid() % 200)]


@pytest.fixture
def free_port():
    import socket
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port
