"""Node-row Linear with the bias gradient from hgk_colsum.

y = x W^T + b and the weight / input gradients stay cuBLAS GEMMs (fp32, TF32 off).  The bias gradient
-- a column sum of the [N, out] output gradient -- is what torch runs as a 128-block reduction for
~20 us at N = 8192; `hgk_colsum` does it in two small launches with a fixed summation order.

The custom autograd node costs more HOST time than the reduction saves, so it is only used where
host time is free: while a CUDA graph is being captured (hydragnn/globalAtt/graphed_tail.py).
Everywhere else `linear_rows` is `F.linear`.  The reference has no counterpart: these are plain
`torch.nn.Linear` calls there (hydragnn/globalAtt/gps.py:58-83)."""
import torch
import torch.nn.functional as F

from .. import _hgk


def colsum(x):
    """[C] = sum over the rows of a [N, C] fp32 CUDA matrix (unit column stride)."""
    if x.stride(1) != 1:
        x = x.contiguous()
    n, c = x.shape
    slabs = _hgk.lib().hgk_colsum_slabs(n)
    out = torch.empty(c, dtype=torch.float32, device=x.device)
    partial = torch.empty((slabs, c), dtype=torch.float32, device=x.device)
    _hgk.call("hgk_colsum", _hgk.ptr(out), _hgk.ptr(partial), _hgk.ptr(x), x.stride(0), n, c,
              cost=(4 * (n * c + 2 * slabs * c + c), 0))
    return out


class _LinearRows(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias):
        ctx.save_for_backward(x, weight)
        return F.linear(x, weight, bias)

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g):
        x, weight = ctx.saved_tensors
        gx = g @ weight if ctx.needs_input_grad[0] else None
        gw = g.t() @ x if ctx.needs_input_grad[1] else None
        gb = colsum(g) if ctx.needs_input_grad[2] else None
        return gx, gw, gb


def linear_rows(x, weight, bias=None):
    if (bias is not None and x.is_cuda and x.dim() == 2 and x.dtype == torch.float32
            and torch.is_grad_enabled() and torch.cuda.is_current_stream_capturing()):
        return _LinearRows.apply(x, weight, bias)
    return F.linear(x, weight, bias)
