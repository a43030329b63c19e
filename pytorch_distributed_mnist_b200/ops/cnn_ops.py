"""CNN layers on the hand-written sm_100a kernels, as autograd Functions.

Pipeline of ``models.CNN.forward`` on CUDA (activations NHWC viewed as ``[B*26*26, C]``)::

    Conv1Fn   direct 3x3 conv + bias + ReLU                      (cnn_kernels.cu, FMA)
    Conv2Fn   implicit-GEMM 3x3 conv, raw accumulators           (gemm_tcgen05.cu, tcgen05/TMA)
    PoolFn    max-pool + bias + ReLU + dropout, NCHW-flat out    (cnn_kernels.cu)
    Fc1Fn     split-K GEMM + bias + ReLU                         (gemm_tcgen05.cu)
    FcOutFn   dropout + Linear(H,10)                             (cnn_kernels.cu, FMA)

Backward runs wgrad/dgrad through the same tcgen05 GEMM (MN-major operands: no transposed
copies).  Internal contract between adjacent Functions: the gradient handed to ``Fc1Fn`` and
``Conv1Fn`` is already multiplied by that layer's ReLU mask (the producer fuses it), which is
idempotent, so each Function is still a correct autograd node on its own.

Parity: the reference has no CNN; the layer list is the PyTorch-examples MNIST net
that BASELINE.json's configs 2-4 name.  Each Function replaces the cuDNN / cuBLAS /
ATen call the stock module would make for that layer (the "stock" arm of
``benchmarks/bench_cnn.py``); numerics are checked against those stock ops in
``tests/test_gpu_cnn.py``.

The module's parameters stay in PyTorch layout (checkpoints are unchanged); conv2's weight is
re-laid out for the GEMM by a tiny kernel each step.
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib
from . import gemm as G

c_void_p, c_int, c_float = ctypes.c_void_p, ctypes.c_int, ctypes.c_float
c_u64 = ctypes.c_uint64


@_lib.optional_decl('pdm_conv1_fwd')
def _d1(fn):
    fn.argtypes = [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p]


@_lib.optional_decl('pdm_conv1_bwd')
def _d2(fn):
    fn.argtypes = [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p]


@_lib.optional_decl('pdm_pool_fwd')
def _d3(fn):
    fn.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_u64, c_void_p]


@_lib.optional_decl('pdm_pool_bwd')
def _d4(fn):
    fn.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_u64, c_void_p]


@_lib.optional_decl('pdm_fcout_fwd')
def _d5(fn):
    fn.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_u64, c_void_p]


@_lib.optional_decl('pdm_fcout_bwd')
def _d6(fn):
    fn.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_float,
                   c_u64, c_void_p]


@_lib.optional_decl('pdm_bias_relu')
def _d7(fn):
    fn.argtypes = [c_void_p, c_void_p, ctypes.c_int64, c_int, c_void_p]


@_lib.optional_decl('pdm_conv_relayout')
def _d8(fn):
    fn.argtypes = [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]


IMG = 26                     # conv1 output is 26x26; conv2's valid output is the 24x24 corner
PIX = IMG * IMG
SM_COUNT = 148


def available():
    return _lib.available() and hasattr(_lib.lib(), 'pdm_gemm_tf32') and hasattr(_lib.lib(), 'pdm_pool_fwd')


def _chk(rc, what):
    _lib.check(rc, what)


def _splits_for(tiles, kblocks):
    want = max(1, (2 * SM_COUNT + tiles - 1) // tiles)
    return max(1, min(want, kblocks))


class Conv1Fn(torch.autograd.Function):

    @staticmethod
    def forward(ctx, x, weight, bias):
        x = x.contiguous()
        B, c1 = x.size(0), weight.size(0)
        out = torch.empty(B * PIX, c1, device=x.device, dtype=torch.float32)
        _chk(_lib.lib().pdm_conv1_fwd(x.data_ptr(), _lib.dtype_code(x), weight.data_ptr(), bias.data_ptr(),
                                      out.data_ptr(), B, c1, _lib.stream_ptr()), 'pdm_conv1_fwd')
        ctx.save_for_backward(x, out)
        ctx.c1 = c1
        return out

    @staticmethod
    def backward(ctx, g):
        x, out = ctx.saved_tensors
        g = (g * (out > 0)).contiguous() if not getattr(g, '_pdm_relu_masked', False) else g.contiguous()
        dw = torch.zeros(ctx.c1, 1, 3, 3, device=x.device)
        db = torch.zeros(ctx.c1, device=x.device)
        _chk(_lib.lib().pdm_conv1_bwd(x.data_ptr(), _lib.dtype_code(x), g.data_ptr(), dw.data_ptr(), db.data_ptr(),
                                      x.size(0), ctx.c1, _lib.stream_ptr()), 'pdm_conv1_bwd')
        return None, dw, db


class Conv2Fn(torch.autograd.Function):

    @staticmethod
    def forward(ctx, a1, weight):
        c2, c1 = weight.size(0), weight.size(1)
        if c1 % 32 or c2 % 32:
            raise ValueError('tcgen05 conv needs channel counts that are multiples of 32')
        lib = _lib.lib()
        need_bwd = ctx.needs_input_grad[0] or ctx.needs_input_grad[1]
        wk = torch.empty(c2, 9 * c1, device=a1.device)
        wd = torch.empty(c1, 9 * c2, device=a1.device) if need_bwd else None
        _chk(lib.pdm_conv_relayout(weight.data_ptr(), wk.data_ptr(), wd.data_ptr() if wd is not None else None,
                                   c2, c1, 0, _lib.stream_ptr()), 'pdm_conv_relayout')
        z = G.gemm(a1, wk, mode=G.MODE_CONV, img_w=IMG, cblocks=c1 // 32, shift_sign=1,
                   M=a1.size(0), N=c2, K=9 * c1)
        ctx.save_for_backward(a1, wd if wd is not None else wk)
        ctx.dims = (c1, c2)
        return z

    @staticmethod
    def backward(ctx, dy):
        a1, wd = ctx.saved_tensors
        c1, c2 = ctx.dims
        dy = dy.contiguous()
        rows = a1.size(0)
        lib = _lib.lib()
        dw = da = None
        if ctx.needs_input_grad[1]:
            tiles = ((c2 + 127) // 128) * ((9 * c1 + 255) // 256)
            g = G.gemm(dy, a1, a_mn_major=True, b_mn_major=True, mode=G.MODE_WGRAD, img_w=IMG, cblocks=c1 // 32,
                       M=c2, N=9 * c1, K=rows, splits=_splits_for(tiles, (rows + 31) // 32))
            dw = torch.zeros(c2, c1, 3, 3, device=dy.device)
            _chk(lib.pdm_conv_relayout(g.data_ptr(), dw.data_ptr(), None, c2, c1, 1, _lib.stream_ptr()),
                 'pdm_conv_relayout')
        if ctx.needs_input_grad[0]:
            # dgrad; the epilogue zeroes positions where conv1's ReLU was inactive (mask = a1 > 0)
            da = G.gemm(dy, wd, mode=G.MODE_CONV, img_w=IMG, cblocks=c2 // 32, shift_sign=-1,
                        M=rows, N=c1, K=9 * c2, mask=a1)
            da._pdm_relu_masked = True
        return da, dw


class PoolFn(torch.autograd.Function):

    @staticmethod
    def forward(ctx, z, bias, p, seed):
        z = z.contiguous()
        C = z.size(1)
        B = z.size(0) // PIX
        out = torch.empty(B, C * 144, device=z.device)
        code = torch.empty(B * C * 144, device=z.device, dtype=torch.uint8)
        _chk(_lib.lib().pdm_pool_fwd(z.data_ptr(), bias.data_ptr(), out.data_ptr(), code.data_ptr(), B, C,
                                     float(p), int(seed), _lib.stream_ptr()), 'pdm_pool_fwd')
        ctx.save_for_backward(code)
        ctx.meta = (B, C, float(p), int(seed))
        return out

    @staticmethod
    def backward(ctx, dp):
        (code,) = ctx.saved_tensors
        B, C, p, seed = ctx.meta
        dp = dp.contiguous()
        dy = torch.empty(B * PIX, C, device=dp.device)
        db = torch.zeros(C, device=dp.device)
        _chk(_lib.lib().pdm_pool_bwd(dp.data_ptr(), code.data_ptr(), dy.data_ptr(), db.data_ptr(), B, C, p, seed,
                                     _lib.stream_ptr()), 'pdm_pool_bwd')
        return dy, db, None, None


class Fc1Fn(torch.autograd.Function):

    @staticmethod
    def forward(ctx, x, weight, bias):
        x = x.contiguous()
        B, K = x.shape
        H = weight.size(0)
        block_n = 128 if H >= 128 else ((H + 31) // 32) * 32
        tiles = ((B + 127) // 128) * ((H + block_n - 1) // block_n)
        splits = _splits_for(tiles, (K + 31) // 32)
        if splits > 1:
            z = torch.zeros(B, H, device=x.device)
            G.gemm(x, weight, out=z, block_n=block_n, splits=splits)
            _chk(_lib.lib().pdm_bias_relu(z.data_ptr(), bias.data_ptr(), B, H, _lib.stream_ptr()), 'pdm_bias_relu')
        else:
            z = G.gemm(x, weight, bias=bias, relu=True, block_n=block_n)
        ctx.save_for_backward(x, weight, z)
        return z

    @staticmethod
    def backward(ctx, g):
        x, weight, h = ctx.saved_tensors
        masked = getattr(g, '_pdm_relu_masked', False)
        db = getattr(g, '_pdm_colsum', None)
        dz = g.contiguous() if masked else (g * (h > 0)).contiguous()
        dx = dw = None
        if ctx.needs_input_grad[1]:
            # dW[H, K] = dz^T . x : both operands MN-major (stored [B][H] and [B][K])
            dw = G.gemm(dz, x, a_mn_major=True, b_mn_major=True, block_n=64)
        if ctx.needs_input_grad[0]:
            # dX[B, K] = dz . W : W stored [H][K] is the MN-major B operand
            dx = G.gemm(dz, weight, b_mn_major=True, block_n=256 if weight.size(1) >= 256 else 64)
        if ctx.needs_input_grad[2] and db is None:
            db = dz.sum(0)
        return dx, dw, db


class FcOutFn(torch.autograd.Function):

    @staticmethod
    def forward(ctx, h, weight, bias, p, seed):
        h = h.contiguous()
        B, H = h.shape
        if weight.size(0) != 10:
            raise ValueError('FcOutFn is specialised for 10 classes')
        logits = torch.empty(B, 10, device=h.device)
        _chk(_lib.lib().pdm_fcout_fwd(h.data_ptr(), weight.data_ptr(), bias.data_ptr(), logits.data_ptr(), B, H,
                                      float(p), int(seed), _lib.stream_ptr()), 'pdm_fcout_fwd')
        ctx.save_for_backward(h, weight)
        ctx.meta = (float(p), int(seed))
        return logits

    @staticmethod
    def backward(ctx, dl):
        h, weight = ctx.saved_tensors
        p, seed = ctx.meta
        B, H = h.shape
        dl = dl.contiguous()
        dz1 = torch.empty(B, H, device=h.device)
        dw = torch.zeros(10, H, device=h.device)
        db = torch.zeros(10, device=h.device)
        db1 = torch.zeros(H, device=h.device)
        _chk(_lib.lib().pdm_fcout_bwd(dl.data_ptr(), h.data_ptr(), weight.data_ptr(), dz1.data_ptr(), dw.data_ptr(),
                                      db.data_ptr(), db1.data_ptr(), B, H, p, seed, _lib.stream_ptr()), 'pdm_fcout_bwd')
        dz1._pdm_relu_masked = True      # (h > 0) already applied: h is fc1's post-ReLU output
        dz1._pdm_colsum = db1
        return dz1, dw, db, None, None


def _seed():
    return int(torch.randint(0, 2 ** 62, (1,)).item())


def cnn_forward(model, x):
    """``models.CNN.forward`` on the native kernels.  ``x``: [B,1,28,28] fp32 or [B,784] uint8."""
    B = x.size(0)
    x2 = x.reshape(B, 784)
    p1 = model.p1 if model.training else 0.0
    p2 = model.p2 if model.training else 0.0
    a1 = Conv1Fn.apply(x2, model.conv1.weight, model.conv1.bias)
    z2 = Conv2Fn.apply(a1, model.conv2.weight)
    pooled = PoolFn.apply(z2, model.conv2.bias, p1, _seed() if p1 > 0 else 0)
    h1 = Fc1Fn.apply(pooled, model.fc1.weight, model.fc1.bias)
    return FcOutFn.apply(h1, model.fc2.weight, model.fc2.bias, p2, _seed() if p2 > 0 else 0)
