"""Python binding of the tcgen05/TMA tf32 GEMM (``csrc/gemm_tcgen05.cu``).

``gemm(a, b, ...)`` computes ``D[M,N] (+)= A[M,K] @ B[N,K]^T`` where each operand is a
row-major fp32 CUDA matrix used as stored (``[MN, K]``, K-major) or transposed
(``[K, MN]``, MN-major) -- no transposed copies are made.  ``conv3x3_*`` are the
implicit-GEMM modes over NHWC activations (see the .cu header for the shifted-row trick).

Parity: stands where the reference stack calls cuBLAS SGEMM (``nn.Linear``) and,
for the CNN extension, cuDNN convolution forward / dgrad / wgrad.
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib


class GemmArgs(ctypes.Structure):
    _fields_ = [('a', ctypes.c_void_p), ('a_rows', ctypes.c_int64), ('a_cols', ctypes.c_int64), ('lda', ctypes.c_int64),
                ('a_mn_major', ctypes.c_int32),
                ('b', ctypes.c_void_p), ('b_rows', ctypes.c_int64), ('b_cols', ctypes.c_int64), ('ldb', ctypes.c_int64),
                ('b_mn_major', ctypes.c_int32),
                ('d', ctypes.c_void_p), ('ldd', ctypes.c_int64),
                ('M', ctypes.c_int32), ('N', ctypes.c_int32), ('K', ctypes.c_int32),
                ('bias', ctypes.c_void_p), ('relu', ctypes.c_int32),
                ('mask', ctypes.c_void_p), ('ld_mask', ctypes.c_int64),
                ('block_n', ctypes.c_int32), ('splits', ctypes.c_int32), ('accumulate', ctypes.c_int32),
                ('mode', ctypes.c_int32), ('img_w', ctypes.c_int32), ('cblocks', ctypes.c_int32),
                ('shift_sign', ctypes.c_int32)]


@_lib.optional_decl('pdm_gemm_tf32')
def _decl(fn):
    fn.argtypes = [ctypes.POINTER(GemmArgs), ctypes.c_void_p]


MODE_GEMM, MODE_CONV, MODE_WGRAD = 0, 1, 2


def _mat(t):
    if t.dim() != 2 or t.dtype != torch.float32 or not t.is_cuda or t.stride(1) != 1:
        raise ValueError('operands must be 2-D fp32 CUDA matrices with unit inner stride')
    return t


def gemm(a, b, out=None, a_mn_major=False, b_mn_major=False, bias=None, relu=False, mask=None,
         block_n=0, splits=1, accumulate=False, mode=MODE_GEMM, img_w=0, cblocks=1, shift_sign=1,
         M=None, N=None, K=None):
    a, b = _mat(a), _mat(b)
    if M is None:
        M = a.size(1) if a_mn_major else a.size(0)
    if K is None:
        K = a.size(0) if a_mn_major else a.size(1)
    if N is None:
        N = b.size(1) if b_mn_major else b.size(0)
    if out is None:
        out = (torch.zeros if (accumulate or splits > 1) else torch.empty)((M, N), device=a.device, dtype=torch.float32)
    g = GemmArgs()
    g.a, g.a_rows, g.a_cols, g.lda, g.a_mn_major = a.data_ptr(), a.size(0), a.size(1), a.stride(0), int(a_mn_major)
    g.b, g.b_rows, g.b_cols, g.ldb, g.b_mn_major = b.data_ptr(), b.size(0), b.size(1), b.stride(0), int(b_mn_major)
    g.d, g.ldd = out.data_ptr(), out.stride(0)
    g.M, g.N, g.K = int(M), int(N), int(K)
    g.bias = bias.data_ptr() if bias is not None else None
    g.relu = int(relu)
    g.mask = mask.data_ptr() if mask is not None else None
    g.ld_mask = mask.stride(0) if mask is not None else 0
    g.block_n, g.splits, g.accumulate = int(block_n), int(splits), int(accumulate)
    g.mode, g.img_w, g.cblocks, g.shift_sign = int(mode), int(img_w), int(cblocks), int(shift_sign)
    _lib.check(_lib.lib().pdm_gemm_tf32(ctypes.byref(g), _lib.stream_ptr()), 'pdm_gemm_tf32')
    return out
