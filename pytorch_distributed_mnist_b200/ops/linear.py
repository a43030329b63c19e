"""``Net.forward`` on CUDA: ``logits = x @ W^T + b`` by the hand-written sm_100a
kernel (``pdm_net_forward``; replaces the cuBLAS ``addmm`` of ref:126).

Only the (784 -> 10) geometry of the reference model has a native kernel; any
other shape, or a non-contiguous weight/bias pair, uses ``F.linear``.  The
backward of this op (used only when someone trains ``Net`` outside the fused
trainer) is two small torch matmuls: the fused step kernel is the training path.
"""
import torch
import torch.nn.functional as F

from . import _lib


def _flat_params(weight, bias):
    """The (W | b) pair as one PDM_NET_FLAT buffer if they already live that way."""
    flat = getattr(weight, '_pdm_flat', None)
    if flat is not None and flat.numel == _lib.NET_FLAT and weight.data_ptr() == flat.params.data_ptr():
        return flat.params
    if weight.data_ptr() + _lib.NET_IN * _lib.NET_OUT * 4 == bias.data_ptr():
        return weight
    return None


class _NetLinear(torch.autograd.Function):

    @staticmethod
    def forward(ctx, x, weight, bias):
        x = x.contiguous()
        params = _flat_params(weight, bias)
        if params is None:
            params = torch.cat([weight.reshape(-1), bias.reshape(-1), weight.new_zeros(2)])
        out = torch.empty(x.size(0), _lib.NET_OUT, device=x.device, dtype=torch.float32)
        _lib.check(_lib.lib().pdm_net_forward(x.data_ptr(), _lib.dtype_code(x), x.size(0), params.data_ptr(),
                                              out.data_ptr(), _lib.stream_ptr()), 'pdm_net_forward')
        ctx.save_for_backward(x, weight)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        x, weight = ctx.saved_tensors
        xf = x if x.dtype == torch.float32 else normalize_u8(x)
        gx = grad_out @ weight if ctx.needs_input_grad[0] and x.dtype == torch.float32 else None
        gw = grad_out.t() @ xf if ctx.needs_input_grad[1] else None
        gb = grad_out.sum(0) if ctx.needs_input_grad[2] else None
        return gx, gw, gb


def normalize_u8(x_u8):
    from ..data import mnist
    return x_u8.to(torch.float32).div_(255.0).sub_(mnist.MEAN).div_(mnist.STD)


def linear(x, weight, bias):
    ok = (x.is_cuda and x.dim() == 2 and x.size(1) == _lib.NET_IN and tuple(weight.shape) == (_lib.NET_OUT, _lib.NET_IN)
          and bias is not None and x.dtype in (torch.float32, torch.uint8) and x.size(0) >= 1)
    if not ok:
        return F.linear(x, weight, bias)
    _lib.lib()   # raises loudly on a GPU box without the native library
    return _NetLinear.apply(x, weight, bias)
