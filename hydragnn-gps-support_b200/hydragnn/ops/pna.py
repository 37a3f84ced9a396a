"""PNA multi-aggregation op (hgk_pna_aggregate_fwd / _bwd).

Replaces PyG PNAConv.propagate + the mean/min/max/std aggregators of
DegreeScalerAggregation as the reference configures them
(hydragnn/models/PNAStack.py:30-53).  The per-edge message
h_e = (P[dst] + Q[src] (+ R_e)) (* M_e) is formed inside the kernel; the optional multiplier M is
PNAPlus's Hadamard product with the projected radial basis (PNAPlusStack.py:275-283).
"""
import torch

from .. import _hgk
from .contract import segment_sum


def _pna_fwd_raw(P, Q, R, topo, M=None):
    N, F = P.shape
    rowptr, perm = topo.csr(1)  # segments = destination (edge_index[1])
    out = torch.empty((N, 4 * F), dtype=torch.float32, device=P.device)
    cmin = torch.empty((N, F), dtype=torch.int32, device=P.device)
    cmax = torch.empty((N, F), dtype=torch.int32, device=P.device)
    E = topo.E
    nbytes = E * (4 * F * (1 + (R is not None) + (M is not None)) + 8) + N * (4 * F + 6 * 4 * F + 4)
    _hgk.call("hgk_pna_aggregate_fwd", _hgk.ptr(out), _hgk.ptr(cmin), _hgk.ptr(cmax), _hgk.ptr(P),
              _hgk.ptr(Q), _hgk.ptr(R), _hgk.ptr(M), _hgk.ptr(rowptr), _hgk.ptr(perm), _hgk.ptr(topo.idx[0]),
              N, E, F, cost=(nbytes, 0))
    return out, cmin, cmax


def _pna_bwd_raw(gout, out, cmin, cmax, P, Q, R, topo, M=None):
    """-> (gh [E,F] = d loss / d (P[dst]+Q[src]+R), gm [E,F] = d loss / d M or None)."""
    N, F = P.shape
    rowptr, perm = topo.csr(1)
    gh = torch.empty((topo.E, F), dtype=torch.float32, device=P.device)
    gm = torch.empty_like(gh) if M is not None else None
    E = topo.E
    nbytes = E * (4 * F * (2 + (R is not None) + 2 * (M is not None)) + 8) + N * (4 * F * 11 + 4)
    _hgk.call("hgk_pna_aggregate_bwd", _hgk.ptr(gh), _hgk.ptr(gm), _hgk.ptr(gout.contiguous()), _hgk.ptr(out),
              _hgk.ptr(cmin), _hgk.ptr(cmax), _hgk.ptr(P), _hgk.ptr(Q), _hgk.ptr(R), _hgk.ptr(M),
              _hgk.ptr(rowptr), _hgk.ptr(perm), _hgk.ptr(topo.idx[0]), N, E, F, cost=(nbytes, 0))
    return gh, gm


class PNAAggregate(torch.autograd.Function):
    @staticmethod
    def forward(ctx, P, Q, R, M, topo):
        P = P.contiguous()
        Q = Q.contiguous()
        R = None if R is None else R.contiguous()
        M = None if M is None else M.contiguous()
        out, cmin, cmax = _pna_fwd_raw(P, Q, R, topo, M)
        ctx.save_for_backward(P, Q, R, M, out, cmin, cmax)
        ctx.topo = topo
        return out

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, gout):
        P, Q, R, M, out, cmin, cmax = ctx.saved_tensors
        topo = ctx.topo
        gh, gm = _pna_bwd_raw(gout, out, cmin, cmax, P, Q, R, topo, M)
        gP = segment_sum(gh, topo, 1) if ctx.needs_input_grad[0] else None
        gQ = segment_sum(gh, topo, 0) if ctx.needs_input_grad[1] else None
        gR = gh if (R is not None and ctx.needs_input_grad[2]) else None
        gM = gm if (M is not None and ctx.needs_input_grad[3]) else None
        return gP, gQ, gR, gM, None


def pna_aggregate(P, Q, R, topo, M=None):
    """[N, 4F] = (mean | min | max | std) over each destination's incoming edges of the message
    (P[dst] + Q[src] + R) * M."""
    return PNAAggregate.apply(P, Q, R, M, topo)
