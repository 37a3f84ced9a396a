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


def _filter_bwd_raw(gw, h, dist, cutoff, w1t):
    E, F = gw.shape
    dz0 = torch.empty_like(gw)
    _hgk.call("hgk_schnet_filter_bwd", _hgk.ptr(dz0), _hgk.ptr(gw), _hgk.ptr(h), _hgk.ptr(dist), float(cutoff),
              _hgk.ptr(w1t), E, F, cost=(E * (12 * F + 4), 2 * E * F * F))
    return dz0


def _wgrad_raw(a, scale_d, cutoff, b=None, gauss=None):
    """(dW [128, KB], db [128]) = (A^T B, colsum A); gauss = (dist, raw, offset, coeff) rebuilds B =
    [gauss(d) | raw] on the fly (KB = 64, the padded layout of the forward kernel)."""
    E, F = a.shape
    if a.stride(1) != 1 or a.stride(0) % 4 or a.data_ptr() % 16:
        a = a.contiguous()
    L = _hgk.lib()
    P = L.hgk_wgrad_tn_partials(E)
    KB = 64 if b is None else b.size(1)
    dw = torch.empty((P, F, KB), dtype=torch.float32, device=a.device)
    db = torch.empty((P, F), dtype=torch.float32, device=a.device)
    if b is None:
        dist, raw, offset, coeff = gauss
        R = 0 if raw is None else raw.size(1)
        args = (_hgk.ptr(None), 0, KB, _hgk.ptr(dist), _hgk.ptr(raw), 0 if raw is None else raw.stride(0), R,
                _hgk.ptr(offset), offset.numel(), float(coeff))
        nbytes = E * (4 * F + 4 + 4 * R)
    else:
        args = (_hgk.ptr(b), b.stride(0), KB, _hgk.ptr(None), _hgk.ptr(None), 0, 0, _hgk.ptr(None), 0, 0.0)
        nbytes = E * 4 * (F + KB)
    _hgk.call("hgk_wgrad_tn", _hgk.ptr(dw), _hgk.ptr(db), _hgk.ptr(a), a.stride(0), _hgk.ptr(scale_d), float(cutoff), *args,
              E, F, cost=(nbytes, 2 * E * F * KB))
    gw = torch.empty((F, KB), dtype=torch.float32, device=a.device)
    gb = torch.empty(F, dtype=torch.float32, device=a.device)
    _hgk.call("hgk_wgrad_reduce", _hgk.ptr(gw), _hgk.ptr(gb), _hgk.ptr(dw), _hgk.ptr(db), P, KB, F,
              cost=(4 * P * F * (KB + 1), 0))        # fixed-order sum of the per-CTA partials, one launch
    return gw, gb


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
        gw = gw.contiguous()
        G = offset.numel()
        GP = (G + 3) // 4 * 4
        dz0 = _filter_bwd_raw(gw, h, dist, ctx.cutoff, w1.t().contiguous())
        gw1, gb1 = _wgrad_raw(gw, dist, ctx.cutoff, b=h)                        # dZ1 = gw * envelope(d)
        gw0p, gb0 = _wgrad_raw(dz0, None, ctx.cutoff, gauss=(dist, raw, offset, ctx.coeff))
        gw0g = gw0p[:, :G]
        gw0r = gw0p[:, GP:GP + raw.size(1)] if raw is not None else None
        return None, None, None, None, None, gw0g, gw0r, gb0, gw1, gb1


def schnet_filter(dist, raw, offset, coeff, cutoff, w0g, w0r, b0, w1, b1):
    return SchnetFilter.apply(dist, raw, offset, coeff, cutoff, w0g, w0r, b0, w1, b1)


# ------------------------------------------------------------ filter network + aggregation ---
def _cfconv_filter_bwd_raw(g, x1, topo, h, dist, cutoff, w1t):
    """dZ0 with the filter gradient gW[e] = g[dst(e)] * x1[src(e)] formed inside the kernel."""
    E, F = h.shape
    dz0 = torch.empty_like(h)
    g_t, ldg = _hgk.rows(g)
    x_t, ldx = _hgk.rows(x1)
    _hgk.call("hgk_cfconv_filter_bwd", _hgk.ptr(dz0), _hgk.ptr(g_t), ldg, _hgk.ptr(topo.idx[1]), _hgk.ptr(x_t), ldx,
              _hgk.ptr(topo.idx[0]), _hgk.ptr(h), _hgk.ptr(dist), float(cutoff), _hgk.ptr(w1t), E, F,
              cost=(E * (8 * F + 12) + 2 * g.size(0) * 4 * F, 2 * E * F * F))
    return dz0


def _cfconv_wgrad_raw(g, x1, topo, dist, cutoff, h):
    """(dW1, db1) = ((gW * envelope)^T h, column sums) with gW formed on the fly."""
    E, F = h.shape
    L = _hgk.lib()
    P = L.hgk_wgrad_tn_partials(E)
    dw = torch.empty((P, F, F), dtype=torch.float32, device=h.device)
    db = torch.empty((P, F), dtype=torch.float32, device=h.device)
    g_t, ldg = _hgk.rows(g)
    x_t, ldx = _hgk.rows(x1)
    _hgk.call("hgk_cfconv_wgrad", _hgk.ptr(dw), _hgk.ptr(db), _hgk.ptr(g_t), ldg, _hgk.ptr(topo.idx[1]), _hgk.ptr(x_t),
              ldx, _hgk.ptr(topo.idx[0]), _hgk.ptr(dist), float(cutoff), _hgk.ptr(h), h.stride(0), E, F,
              cost=(E * (4 * F + 12) + 2 * g.size(0) * 4 * F, 2 * E * F * F))
    gw = torch.empty((F, F), dtype=torch.float32, device=h.device)
    gb = torch.empty(F, dtype=torch.float32, device=h.device)
    _hgk.call("hgk_wgrad_reduce", _hgk.ptr(gw), _hgk.ptr(gb), _hgk.ptr(dw), _hgk.ptr(db), P, F, F,
              cost=(4 * P * F * (F + 1), 0))
    return gw, gb


# HYDRAGNN_CFCONV_GATHER=1: form the filter gradient inside the dZ0 / dW1 kernels instead of materialising it
import os as _os
GATHER_IN_KERNEL = _os.environ.get("HYDRAGNN_CFCONV_GATHER", "0") == "1"


class CFConvFused(torch.autograd.Function):
    """out[t] = sum_{e: dst(e) = t} x1[src(e)] * W_e with W = filter network(d, raw) * envelope
    (reference SCFStack.py:268-293: the `nn` Sequential, the cosine cutoff and PyG's propagate) as ONE
    autograd node (one python-level backward instead of five).  Forward: the fused filter kernel, then
    the CSR aggregation.  Backward, two variants of the filter gradient gW[e] = g[dst(e)] * x1[src(e)]:
      * default: hgk_contract_edge_w writes it once, the dZ0 and dW1 kernels stream it;
      * GATHER_IN_KERNEL: never written -- hgk_cfconv_filter_bwd / hgk_cfconv_wgrad form it from the
        two node tables (4 MB each, L2-resident) while they stage their operands: one [E, F] write and
        two [E, F] reads less per layer (730 MB at config 5), but measured SLOWER on B200 (dZ0 0.21 ->
        0.26 ms, dW1 0.15 -> 0.23 ms per layer at E = 474k): the two dependent gathers per row sit on
        the staging path that already limits those kernels (LSU, not HBM).
    First derivatives only (same condition as `schnet_filter`)."""

    @staticmethod
    def forward(ctx, x1, topo, dist, raw, offset, coeff, cutoff, w0g, w0r, b0, w1, b1):
        from .contract import _node_raw

        F, G = w0g.shape
        GP = (G + 3) // 4 * 4
        w0p = torch.zeros((F, 64), dtype=torch.float32, device=dist.device)
        w0p[:, :G] = w0g
        if raw is not None:
            raw = raw.contiguous()
            w0p[:, GP:GP + raw.size(1)] = w0r
        dist = dist.contiguous()
        x1 = x1.contiguous()
        w, h = _filter_fwd_raw(dist, raw, offset, coeff, cutoff, w0p, b0.contiguous(), w1.contiguous(), b1.contiguous())
        out = _node_raw(topo, 1, x1, 0, w, None, 1, 1, F)
        ctx.save_for_backward(x1, dist, raw, offset, h, w, w1)
        ctx.topo, ctx.coeff, ctx.cutoff = topo, coeff, cutoff
        return out

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g):
        from .contract import _node_raw

        x1, dist, raw, offset, h, w, w1 = ctx.saved_tensors
        topo = ctx.topo
        F = h.size(1)
        G = offset.numel()
        GP = (G + 3) // 4 * 4
        g = g.contiguous()
        dx1 = _node_raw(topo, 0, g, 1, w, None, 1, 1, F) if ctx.needs_input_grad[0] else None
        if GATHER_IN_KERNEL:
            dz0 = _cfconv_filter_bwd_raw(g, x1, topo, h, dist, ctx.cutoff, w1.t().contiguous())
            gw1, gb1 = _cfconv_wgrad_raw(g, x1, topo, dist, ctx.cutoff, h)
        else:
            from .contract import _edge_w_raw
            gw = _edge_w_raw(topo, g, 1, x1, 0, None, 1, F)
            dz0 = _filter_bwd_raw(gw, h, dist, ctx.cutoff, w1.t().contiguous())
            gw1, gb1 = _wgrad_raw(gw, dist, ctx.cutoff, b=h)
        gw0p, gb0 = _wgrad_raw(dz0, None, ctx.cutoff, gauss=(dist, raw, offset, ctx.coeff))
        gw0g = gw0p[:, :G]
        gw0r = gw0p[:, GP:GP + raw.size(1)] if raw is not None else None
        return dx1, None, None, None, None, None, None, gw0g, gw0r, gb0, gw1, gb1


def cfconv_fused(x1, topo, dist, raw, offset, coeff, cutoff, w0g, w0r, b0, w1, b1):
    return CFConvFused.apply(x1, topo, dist, raw, offset, coeff, cutoff, w0g, w0r, b0, w1, b1)
