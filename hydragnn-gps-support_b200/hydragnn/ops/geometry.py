"""Edge geometry as one differentiable operator with hand-written first and second derivatives
(hgk_edge_geom_fwd / _bwd / _bwd2).

Reference: get_edge_vectors_and_lengths (hydragnn/utils/model/operations.py:21-36) -- two index_selects,
a subtraction, the shift addition, a vector norm and a division, differentiated by autograd; under the
force loss (Base.py:614-620, create_graph=True) the backward of that chain is itself differentiated,
which costs a few dozen elementwise launches per call.  Here the forward is one kernel, the
vector-Jacobian product t = d loss / d vec one kernel (its scatter to the two endpoints: two segmented
sums of the contraction family, deterministic), and the derivative of that map one more kernel.
"""
import torch

from .. import _hgk
from .contract import segment_sum


def _fwd_raw(pos, src, dst, shifts, eps, want_unit):
    E = src.numel()
    vec = torch.empty((E, 3), dtype=torch.float32, device=pos.device)
    length = torch.empty(E, dtype=torch.float32, device=pos.device)
    unit = torch.empty((E, 3), dtype=torch.float32, device=pos.device) if want_unit else None
    _hgk.call("hgk_edge_geom_fwd", _hgk.ptr(vec), _hgk.ptr(length), _hgk.ptr(unit), _hgk.ptr(pos), _hgk.ptr(src),
              _hgk.ptr(dst), _hgk.ptr(shifts), float(eps), E, cost=(E * (8 + 24 + 16 + (12 if shifts is not None else 0)
                                                                        + (12 if want_unit else 0)), 0))
    return vec, length, unit


def _bwd_raw(g_vec, g_len, g_unit, vec, length, eps):
    E = vec.size(0)
    t = torch.empty_like(vec)
    _hgk.call("hgk_edge_geom_bwd", _hgk.ptr(t), _hgk.ptr(g_vec), _hgk.ptr(g_len), _hgk.ptr(g_unit), _hgk.ptr(vec),
              _hgk.ptr(length), float(eps), E, cost=(E * 56, 0))
    return t


def _bwd2_raw(h, g_len, g_unit, vec, length, eps):
    E = vec.size(0)
    d_glen = torch.empty(E, dtype=torch.float32, device=vec.device)
    d_gunit = torch.empty_like(vec)
    d_vec = torch.empty_like(vec)
    _hgk.call("hgk_edge_geom_bwd2", _hgk.ptr(d_glen), _hgk.ptr(d_gunit), _hgk.ptr(d_vec), _hgk.ptr(h), _hgk.ptr(g_len),
              _hgk.ptr(g_unit), _hgk.ptr(vec), _hgk.ptr(length), float(eps), E, cost=(E * 84, 0))
    return d_glen, d_gunit, d_vec


def _c(t):
    return None if t is None else t.contiguous()


class _EdgeGeomVjp(torch.autograd.Function):
    """t = d loss / d vec from the gradients of (vec, len, unit); differentiable once more (bwd2)."""

    @staticmethod
    def forward(ctx, g_vec, g_len, g_unit, vec, length, eps):
        g_vec, g_len, g_unit = _c(g_vec), _c(g_len), _c(g_unit)
        ctx.save_for_backward(g_len, g_unit, vec, length)
        ctx.eps = eps
        return _bwd_raw(g_vec, g_len, g_unit, vec, length, eps)

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, h):
        g_len, g_unit, vec, length = ctx.saved_tensors
        h = h.contiguous()
        d_glen, d_gunit, d_vec = _bwd2_raw(h, g_len, g_unit, vec, length, ctx.eps)
        return (h if ctx.needs_input_grad[0] else None, d_glen if ctx.needs_input_grad[1] else None,
                d_gunit if ctx.needs_input_grad[2] else None, d_vec if ctx.needs_input_grad[3] else None, None, None)


class EdgeGeom(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pos, shifts, topo, eps, want_unit):
        pos = pos.contiguous()
        shifts = _c(shifts)
        vec, length, unit = _fwd_raw(pos, topo.idx[0], topo.idx[1], shifts, eps, want_unit)
        ctx.save_for_backward(vec, length)
        ctx.topo, ctx.eps, ctx.has_shifts, ctx.want_unit = topo, eps, shifts is not None, want_unit
        if want_unit:
            return vec, length, unit
        return vec, length

    @staticmethod
    def backward(ctx, g_vec, g_len, g_unit=None):
        vec, length = ctx.saved_tensors
        # `vec` is an output of this node: gradients that reach it through the differentiated
        # backward (second order) come back here as g_vec; `length` = |vec| is handled inside the
        # formulas, so it enters as a constant
        t = _EdgeGeomVjp.apply(g_vec, g_len, g_unit if ctx.want_unit else None, vec, length.detach(), ctx.eps)
        d_pos = None
        if ctx.needs_input_grad[0]:
            d_pos = segment_sum(t, ctx.topo, 1) - segment_sum(t, ctx.topo, 0)
        d_shifts = t if (ctx.has_shifts and ctx.needs_input_grad[1]) else None
        return d_pos, d_shifts, None, None, None


def edge_geometry(pos, topo, shifts=None, eps=1e-9, want_unit=False):
    """-> (vec [E,3], len [E], unit [E,3] or None) for the edges of `topo` (side 0 = sender, 1 = receiver)."""
    out = EdgeGeom.apply(pos, shifts, topo, float(eps), bool(want_unit))
    return out if want_unit else (out[0], out[1], None)
