"""The allocation-free entry point ``pdm_net_train_step_static`` against ``pdm_net_train_step``.

Some hosts of the library cannot hand the fused step its scratch buffer because they may not
allocate device memory where the step runs (a process that pre-reserved all HBM for its own pool
and forbids later cudaMalloc, code running under another allocator's lock, an out-of-memory
recovery path).  For them the library keeps a fixed scratch array inside the loaded module.
This test drives both entry points through the C ABI (ctypes on the shared library) with
identical inputs and state, two calls in a row, and requires the same parameters, optimizer
state, gradient and per-step outputs; it also checks the size limit and that the second call
leaves the amount of free device memory untouched.

The file name sorts last on purpose: with ``pytest -x`` every other test has already run if
this one fails.  It passes alone in a fresh process and after the rest of the suite.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _args(lib_mod, x, y, params, m, v, grad, step_out, metrics, step, scratch=None, sync=None, ctas=0):
    a = lib_mod.NetStepArgs()
    a.x, a.x_dtype = x.data_ptr(), lib_mod.PDM_U8
    a.y, a.y_dtype = y.data_ptr(), lib_mod.PDM_U8
    a.batch, a.ctas = x.size(0), ctas
    a.params, a.m, a.v = params.data_ptr(), m.data_ptr(), v.data_ptr()
    a.grad_out = grad.data_ptr()
    a.scratch = scratch.data_ptr() if scratch is not None else None
    a.sync = sync.data_ptr() if sync is not None else None
    a.metrics = metrics.data_ptr()
    a.step_out = step_out.data_ptr()
    a.hyper = lib_mod.Hyper(opt=lib_mod.PDM_OPT_ADAM, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8,
                            weight_decay=0.0, step=step)
    return a


def _state(dev, seed):
    g = torch.Generator().manual_seed(seed)
    params = (torch.randn(7852, generator=g) * 0.05).to(dev)
    params[7850:] = 0
    return {'params': params, 'm': torch.zeros(7852, device=dev), 'v': torch.zeros(7852, device=dev),
            'grad': torch.zeros(7852, device=dev), 'step_out': torch.zeros(2, device=dev),
            'metrics': torch.zeros(4, dtype=torch.float64, device=dev)}


def test_static_scratch_entry_point_matches_regular_one():
    from pytorch_distributed_mnist_b200.ops import _lib
    lib = _lib.lib()
    assert hasattr(lib, 'pdm_net_train_step_static')
    dev = torch.device('cuda', 0)
    torch.cuda.set_device(dev)
    stream = _lib.stream_ptr()
    ref, new = _state(dev, 11), _state(dev, 11)
    scratch = torch.empty(_lib.NET_MAX_CTAS * _lib.NET_SCRATCH_STRIDE, device=dev)
    sync = torch.zeros(64, dtype=torch.int32, device=dev)
    free_before_second = None
    for step, batch in ((1, 512), (2, 200)):                       # two calls in a row, second one ragged
        g = torch.Generator().manual_seed(100 + step)
        x = torch.randint(0, 256, (batch, 784), generator=g, dtype=torch.uint8).to(dev)
        y = torch.randint(0, 10, (batch,), generator=g, dtype=torch.uint8).to(dev)
        a = _args(_lib, x, y, ref['params'], ref['m'], ref['v'], ref['grad'], ref['step_out'], ref['metrics'], step,
                  scratch=scratch, sync=sync)
        _lib.check(lib.pdm_net_train_step(ctypes.byref(a), stream), 'pdm_net_train_step')
        torch.cuda.synchronize()
        if step == 2:
            free_before_second = torch.cuda.mem_get_info(dev)[0]
        b = _args(_lib, x, y, new['params'], new['m'], new['v'], new['grad'], new['step_out'], new['metrics'], step)
        _lib.check(lib.pdm_net_train_step_static(ctypes.byref(b), stream), 'pdm_net_train_step_static')
        torch.cuda.synchronize()
        if step == 2:
            assert torch.cuda.mem_get_info(dev)[0] == free_before_second, 'the call changed free device memory'
        for k in ('params', 'm', 'v', 'grad', 'step_out'):
            torch.testing.assert_close(new[k], ref[k], rtol=1e-5, atol=1e-6, msg=lambda m, k=k: '{} (step {}): {}'.format(k, step, m))
        torch.testing.assert_close(new['metrics'], ref['metrics'], rtol=1e-9, atol=1e-9)
    assert not torch.equal(ref['params'], _state(dev, 11)['params'])   # the steps really updated something
    assert int(new['metrics'][2].item()) == 712


def test_static_scratch_entry_point_rejects_oversize_requests():
    from pytorch_distributed_mnist_b200.ops import _lib
    lib = _lib.lib()
    dev = torch.device('cuda', 0)
    torch.cuda.set_device(dev)
    st = _state(dev, 3)
    before = st['params'].clone()
    x = torch.zeros(64, 784, dtype=torch.uint8, device=dev)
    y = torch.zeros(64, dtype=torch.uint8, device=dev)
    a = _args(_lib, x, y, st['params'], st['m'], st['v'], st['grad'], st['step_out'], st['metrics'], 1,
              ctas=_lib.NET_STATIC_MAX_CTAS + 1)
    rc = lib.pdm_net_train_step_static(ctypes.byref(a), _lib.stream_ptr())
    assert rc == -6                                               # PDM_ERR_LIMIT
    msg = lib.pdm_last_error().decode()
    assert 'at most 64' in msg and 'PDM_NET_STATIC_MAX_CTAS' in msg
    torch.cuda.synchronize()
    assert torch.equal(st['params'], before)                      # nothing was launched
    # the largest allowed request works, and so does a batch far larger than 64 tiles
    a = _args(_lib, x, y, st['params'], st['m'], st['v'], st['grad'], st['step_out'], st['metrics'], 1,
              ctas=_lib.NET_STATIC_MAX_CTAS)
    _lib.check(lib.pdm_net_train_step_static(ctypes.byref(a), _lib.stream_ptr()), 'pdm_net_train_step_static')
    xb = torch.randint(0, 256, (5000, 784), dtype=torch.uint8, device=dev)
    yb = torch.randint(0, 10, (5000,), dtype=torch.uint8, device=dev)
    a = _args(_lib, xb, yb, st['params'], st['m'], st['v'], st['grad'], st['step_out'], st['metrics'], 2)
    _lib.check(lib.pdm_net_train_step_static(ctypes.byref(a), _lib.stream_ptr()), 'pdm_net_train_step_static')
    torch.cuda.synchronize()
    assert torch.isfinite(st['params']).all() and not torch.equal(st['params'], before)
