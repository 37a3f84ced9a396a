"""Fused SchNet filter network (hgk_schnet_filter_fwd) as an autograd Function.

W_e = Linear1(ssp(Linear0([gauss(d_e) | raw_e]))) * C(d_e)   -- reference SCFStack.py:259-272.
Forward: one tcgen05 kernel that never materialises the gaussian expansion, the concatenation or
the first Linear's output.  Backward (first derivative only; the fused path is taken when neither
the distances nor the raw attribute require grad, i.e. no force loss): recomputed from the saved
hidden activations.
"""
import math

import torch

from .. import _hgk

_LN2 = math.log(2.0)
SUPPORTED = dict(F=128, G=50, R=(0, 8))


def fused_filter_ok(dist, raw, num_filters, num_gaussians):
    if not (dist.is_cuda and dist.dtype == torch.float32 and not dist.requires_grad):
        return False
    if raw is not None and (raw.requires_grad or raw.dtype != torch.float32):
        return False
    r = 0 if raw is None else raw.size(1)
    return num_filters == SUPPORTED["F"] and num_gaussians == SUPPORTED["G"] and r in SUPPORTED["R"]


def _filter_fwd_raw(dist, raw, offset, coeff, cutoff, w0p, b0, w1, b1):
    E, F = dist.numel(), w1.size(0)
    w = torch.empty((E, F), dtype=torch.float32, device=dist.device)
    h = torch.empty((E, F), dtype=torch.float32, device=dist.device)
    R = 0 if raw is None else raw.size(1)
    ldr = 0 if raw is None else raw.stride(0)
    G = offset.numel()
    flops = 2 * E * F * (G + R + F)
    _hgk.call("hgk_schnet_filter_fwd", _hgk.ptr(w), _hgk.ptr(h), _hgk.ptr(dist), _hgk.ptr(raw), ldr, R,
              _hgk.ptr(offset), G, float(coeff), float(cutoff), _hgk.ptr(w0p), _hgk.ptr(b0), _hgk.ptr(w1),
              _hgk.ptr(b1), E, F, cost=(E * (4 + 4 * R + 8 * F), flops))
    return w, h


class SchnetFilter(torch.autograd.Function):
    @staticmethod
    def forward(ctx, dist, raw, offset, coeff, cutoff, w0g, w0r, b0, w1, b1):
        F, G = w0g.shape
        GP = (G + 3) // 4 * 4
        w0p = torch.zeros((F, 64), dtype=torch.float32, device=dist.device)
        w0p[:, :G] = w0g
        if raw is not None:
            raw = raw.contiguous()
            w0p[:, GP:GP + raw.size(1)] = w0r
        dist = dist.contiguous()
        w, h = _filter_fwd_raw(dist, raw, offset, coeff, cutoff, w0p, b0.contiguous(), w1.contiguous(),
                               b1.contiguous())
        ctx.save_for_backward(dist, raw, offset, h, w1)
        ctx.coeff, ctx.cutoff = coeff, cutoff
        return w

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, gw):
        dist, raw, offset, h, w1 = ctx.saved_tensors
        env = 0.5 * (torch.cos(dist * math.pi / ctx.cutoff) + 1.0)
        dz1 = gw * env.view(-1, 1)
        gb1 = dz1.sum(0)
        gw1 = dz1.t() @ h
        dz0 = (dz1 @ w1) * (1.0 - torch.exp(-(h + _LN2)))          # ssp'(z) = sigmoid(z) = 1 - exp(-softplus(z))
        gb0 = dz0.sum(0)
        delta = dist.view(-1, 1) - offset.view(1, -1)
        gw0g = dz0.t() @ torch.exp(ctx.coeff * delta * delta)
        gw0r = dz0.t() @ raw if raw is not None else None
        return None, None, None, None, None, gw0g, gw0r, gb0, gw1, gb1


def schnet_filter(dist, raw, offset, coeff, cutoff, w0g, w0r, b0, w1, b1):
    return SchnetFilter.apply(dist, raw, offset, coeff, cutoff, w0g, w0r, b0, w1, b1)
