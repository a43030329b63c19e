"""Fused cross-entropy for the generic (autograd) path.

``F.cross_entropy`` (ref:88, ref:111) plus ``Accuracy.update``'s
argmax/eq/sum (ref:61-62) in one kernel: mean loss, dlogits and the top-1 /
loss-sum accumulators (``pdm_cross_entropy``).  dlogits is
produced in the forward pass, so backward is a scalar multiply.
"""
import torch
import torch.nn.functional as F

from . import _lib


class _FusedCE(torch.autograd.Function):

    @staticmethod
    def forward(ctx, logits, target, metrics):
        logits = logits.contiguous()
        target = target.contiguous()
        loss = torch.empty((), device=logits.device, dtype=torch.float32)
        need_grad = ctx.needs_input_grad[0]
        dlogits = torch.empty_like(logits) if need_grad else None
        _lib.check(_lib.lib().pdm_cross_entropy(
            logits.data_ptr(), target.data_ptr(), _lib.dtype_code(target), logits.size(0), logits.size(1),
            loss.data_ptr(), dlogits.data_ptr() if need_grad else None,
            metrics.data_ptr() if metrics is not None else None, _lib.stream_ptr()), 'pdm_cross_entropy')
        if need_grad:
            ctx.save_for_backward(dlogits)
        return loss

    @staticmethod
    def backward(ctx, grad_out):
        (dlogits,) = ctx.saved_tensors
        return dlogits * grad_out, None, None


def cross_entropy(logits, target, metrics=None):
    """Mean cross-entropy.  ``metrics`` (optional float64[4] device tensor) is
    incremented by {loss*B, correct, B, 0} in the same kernel."""
    if (logits.is_cuda and logits.dtype == torch.float32 and logits.dim() == 2
            and target.dtype in (torch.int64, torch.int32, torch.uint8)):
        _lib.lib()
        return _FusedCE.apply(logits, target, metrics)
    loss = F.cross_entropy(logits, target.long())
    if metrics is not None:
        with torch.no_grad():
            n = logits.size(0)
            metrics[0] += loss.detach().double() * n
            metrics[1] += (logits.argmax(dim=1) == target).sum().double()
            metrics[2] += n
    return loss
