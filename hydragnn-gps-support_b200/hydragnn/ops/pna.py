"""PNA multi-aggregation op (hgk_pna_aggregate_fwd / _bwd).

Replaces PyG PNAConv.propagate + the mean/min/max/std aggregators of
DegreeScalerAggregation as the reference configures them
(hydragnn/models/PNAStack.py:30-53).  The per-edge message
h_e = (P[dst] + Q[src] (+ R_e)) (* M_e) is formed inside the kernel; the optional multiplier M is
PNAPlus's Hadamard product with the projected radial basis (PNAPlusStack.py:275-283).
"""
import math

import torch

from .. import _hgk
from .contract import segment_sum


def _pna_fwd_raw(P, Q, R, topo, M=None, side=1):
    """Aggregate over the edges grouped by endpoint `side` (1 = edge_index[1], PyG's target; 0 =
    edge_index[0], where the PaiNN-style stacks aggregate); P rows are taken at that endpoint, Q rows
    at the other one."""
    first = next(t for t in (P, Q, R) if t is not None)
    N, F = topo.num_rows[side], first.size(1)
    rowptr, perm = topo.csr(side)
    dev = first.device
    out = torch.empty((N, 4 * F), dtype=torch.float32, device=dev)
    cmin = torch.empty((N, F), dtype=torch.int32, device=dev)
    cmax = torch.empty((N, F), dtype=torch.int32, device=dev)
    E = topo.E
    terms = (Q is not None) + (R is not None) + (M is not None)
    nbytes = E * (4 * F * terms + 8) + N * (4 * F * (P is not None) + 6 * 4 * F + 4)
    _hgk.call("hgk_pna_aggregate_fwd", _hgk.ptr(out), _hgk.ptr(cmin), _hgk.ptr(cmax), _hgk.ptr(P),
              _hgk.ptr(Q), _hgk.ptr(R), _hgk.ptr(M), _hgk.ptr(rowptr), _hgk.ptr(perm), _hgk.ptr(topo.idx[1 - side]),
              N, E, F, cost=(nbytes, 0))
    return out, cmin, cmax


def _pna_bwd_raw(gout, out, cmin, cmax, P, Q, R, topo, M=None, side=1):
    """-> (gh [E,F] = d loss / d (P+Q+R), gm [E,F] = d loss / d M or None)."""
    N, F = out.size(0), out.size(1) // 4
    rowptr, perm = topo.csr(side)
    gh = torch.empty((topo.E, F), dtype=torch.float32, device=out.device)
    gm = torch.empty_like(gh) if M is not None else None
    E = topo.E
    terms = 1 + (Q is not None) + (R is not None) + 2 * (M is not None)
    nbytes = E * (4 * F * terms + 8) + N * (4 * F * 11 + 4)
    _hgk.call("hgk_pna_aggregate_bwd", _hgk.ptr(gh), _hgk.ptr(gm), _hgk.ptr(gout.contiguous()), _hgk.ptr(out),
              _hgk.ptr(cmin), _hgk.ptr(cmax), _hgk.ptr(P), _hgk.ptr(Q), _hgk.ptr(R), _hgk.ptr(M),
              _hgk.ptr(rowptr), _hgk.ptr(perm), _hgk.ptr(topo.idx[1 - side]), N, E, F, cost=(nbytes, 0))
    return gh, gm


def _aggregate_torch(P, Q, R, M, topo, side):
    """The same four aggregations from differentiable torch ops (scatter_reduce).  Only used to
    differentiate the backward pass itself (create_graph=True: force losses, e.g. the reference's
    physics-informed PNAEq configuration) -- the kernels provide first derivatives only."""
    idx = topo.idx[side].long()
    N = topo.num_rows[side]
    h = None
    for t, gather in ((P, idx), (Q, topo.idx[1 - side].long()), (R, None)):
        if t is None:
            continue
        term = t if gather is None else t.index_select(0, gather)
        h = term if h is None else h + term
    if M is not None:
        h = h * M
    F = h.size(1)
    cnt = torch.zeros(N, dtype=h.dtype, device=h.device).index_add_(0, idx, torch.ones_like(idx, dtype=h.dtype))
    safe = cnt.clamp(min=1).unsqueeze(1)
    mean = torch.zeros((N, F), dtype=h.dtype, device=h.device).index_add(0, idx, h) / safe
    mean2 = torch.zeros((N, F), dtype=h.dtype, device=h.device).index_add(0, idx, h * h) / safe
    expand = idx.unsqueeze(1).expand_as(h)
    mn = torch.zeros((N, F), dtype=h.dtype, device=h.device).scatter_reduce(0, expand, h, "amin", include_self=False)
    mx = torch.zeros((N, F), dtype=h.dtype, device=h.device).scatter_reduce(0, expand, h, "amax", include_self=False)
    sd = (mean2 - mean * mean).clamp(min=1e-5).sqrt()
    sd = sd.masked_fill(sd <= math.sqrt(1e-5), 0.0) * (cnt > 0).unsqueeze(1)
    return torch.cat([mean, mn, mx, sd], dim=1)


class PNAAggregate(torch.autograd.Function):
    @staticmethod
    def forward(ctx, P, Q, R, M, topo, side):
        out, cmin, cmax = _pna_fwd_raw(P, Q, R, topo, M, side)
        ctx.save_for_backward(P, Q, R, M, out, cmin, cmax)
        ctx.topo, ctx.side = topo, side
        return out

    @staticmethod
    def backward(ctx, gout):
        P, Q, R, M, out, cmin, cmax = ctx.saved_tensors
        topo, side = ctx.topo, ctx.side
        if torch.is_grad_enabled():        # create_graph=True: stay differentiable
            inputs = [t for t in (P, Q, R, M) if t is not None and t.requires_grad]
            with torch.enable_grad():
                grads = torch.autograd.grad(_aggregate_torch(P, Q, R, M, topo, side), inputs, gout,
                                            create_graph=True, allow_unused=True)
            it = iter(grads)
            return tuple(next(it) if (t is not None and t.requires_grad) else None for t in (P, Q, R, M)) + (None, None)
        gh, gm = _pna_bwd_raw(gout, out, cmin, cmax, P, Q, R, topo, M, side)
        gP = segment_sum(gh, topo, side) if (P is not None and ctx.needs_input_grad[0]) else None
        gQ = segment_sum(gh, topo, 1 - side) if (Q is not None and ctx.needs_input_grad[1]) else None
        gR = gh if (R is not None and ctx.needs_input_grad[2]) else None
        gM = gm if (M is not None and ctx.needs_input_grad[3]) else None
        return gP, gQ, gR, gM, None, None


def pna_aggregate(P, Q, R, topo, M=None, side=1):
    """[N, 4F] = (mean | min | max | std) over the edges of each node at endpoint `side` of the
    message (P[that node] + Q[other endpoint] + R) * M; P, Q, R (not all) and M may be None."""
    # contiguous copies are made HERE, under autograd, so that the tensors the Function saves are its
    # differentiable inputs (the create_graph path recomputes from them)
    P, Q, R, M = (None if t is None else t.contiguous() for t in (P, Q, R, M))
    return PNAAggregate.apply(P, Q, R, M, topo, side)
