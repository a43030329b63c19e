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

# Numerical mode of every GEMM-shaped layer.
#   'tf32'    one tcgen05 kind::tf32 pass: operands carry 10 explicit mantissa bits, fp32 accumulation
#             (the precision class of cuDNN/cuBLAS with TF32 allowed).  Default, BASELINE.json's choice.
#   'fp32x3'  3xTF32: A = A_hi + A_lo, B = B_hi + B_lo (hi = the 19 bits the tensor core reads, lo = the
#             exact remainder) and D = A_lo.B_hi + A_hi.B_lo + A_hi.B_hi on the same kernel; the dropped
#             A_lo.B_lo term is ~2^-22 relative: fp32-faithful, three passes.
PRECISIONS = ('tf32', 'fp32x3')
_precision = 'tf32'


def set_precision(mode):
    """Select 'tf32' (default) or 'fp32x3' for all GEMM-shaped layers; returns the previous mode."""
    global _precision
    if mode not in PRECISIONS:
        raise ValueError('precision must be one of {}'.format(PRECISIONS))
    prev, _precision = _precision, mode
    return prev


def get_precision():
    return _precision


def _split_tf32(t):
    """hi = t with the 13 mantissa bits a tf32 operand ignores cleared, lo = t - hi (exact in fp32)."""
    hi = (t.view(torch.int32) & -8192).view(torch.float32)
    return hi, t - hi


def _mat(t):
    if t.dim() != 2 or t.dtype != torch.float32 or not t.is_cuda or t.stride(1) != 1:
        raise ValueError('operands must be 2-D fp32 CUDA matrices with unit inner stride')
    return t


def gemm(a, b, out=None, a_mn_major=False, b_mn_major=False, bias=None, relu=False, mask=None,
         block_n=0, splits=1, accumulate=False, mode=MODE_GEMM, img_w=0, cblocks=1, shift_sign=1,
         M=None, N=None, K=None, precision=None):
    a, b = _mat(a), _mat(b)
    if (precision or _precision) == 'fp32x3':
        return _gemm_fp32x3(a, b, out, a_mn_major, b_mn_major, bias, relu, mask, block_n, splits, accumulate, mode, img_w,
                            cblocks, shift_sign, M, N, K)
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


def _gemm_fp32x3(a, b, out, a_mn_major, b_mn_major, bias, relu, mask, block_n, splits, accumulate, mode, img_w, cblocks,
                 shift_sign, M, N, K):
    """Three accumulating tf32 passes over split operands, then the epilogue in fp32 (the fused epilogue of
    the kernel cannot be combined with accumulation)."""
    a_hi, a_lo = _split_tf32(a)
    b_hi, b_lo = _split_tf32(b)
    kw = dict(a_mn_major=a_mn_major, b_mn_major=b_mn_major, block_n=block_n, splits=splits, mode=mode, img_w=img_w,
              cblocks=cblocks, shift_sign=shift_sign, M=M, N=N, K=K, precision='tf32')
    acc = gemm(a_lo, b_hi, **kw)                       # small terms first
    if splits <= 1:
        gemm(a_hi, b_lo, out=acc, accumulate=True, **kw)
        gemm(a_hi, b_hi, out=acc, accumulate=True, **kw)
    else:                                              # split-K already accumulates with atomics into `out`
        gemm(a_hi, b_lo, out=acc, **kw)
        gemm(a_hi, b_hi, out=acc, **kw)
    if bias is not None:
        acc = acc + bias
    if relu:
        acc = torch.relu_(acc)
    if mask is not None:
        acc = acc * (mask > 0)
    if out is not None:
        if accumulate:
            out += acc
        else:
            out.copy_(acc)
        return out
    return acc
