"""Autograd wrappers of the contraction family (hgk_contract_node / _edge_w / _edge_u).

Every gather / multiply / segmented-sum of the reference's SchNet, EGNN and
PAINN convolutions, the mean/sum pooling of Base.forward and the energy sum of
Base.energy_force_loss is one of the three partial derivatives of

    Phi(G, X, W, U) = sum_e sum_c sum_f G[i_e,c,f] X[j_e,c,f] W[e,f] U[e,c]

and the backward of each of the three is made of the other two, so the
Functions below are differentiable to any order.  That is what
`torch.autograd.grad(..., create_graph=True)` in the force loss
(reference hydragnn/models/Base.py:614-620) needs.

Shapes: node operands are [rows, F] (one channel) or [rows, C, F]; W is
[E, F]; U is [E, C].  C is 1 or 3.  `None` means "all ones".
"""
import torch

from .. import _hgk


def _channels(t):
    return 1 if t.dim() == 2 else t.size(1)


def _node_raw(topo, side_out, A, side_a, W, U, C, CO, F):
    N = topo.num_rows[side_out]
    rowptr, perm = topo.csr(side_out)
    dev = rowptr.device
    out = torch.empty((N, F) if CO == 1 else (N, CO, F), dtype=torch.float32, device=dev)
    a_t, lda, ca = None, 0, 1
    if A is not None:
        a_t, lda = _hgk.rows(A)
        ca = _channels(A)
    w_t, ldw = (None, 0) if W is None else _hgk.rows(W)
    u_t = None if U is None else U.contiguous()
    gidx = topo.idx[side_a] if A is not None else None
    E = topo.E
    # algorithmic bytes: gathered A rows + W rows + U + indices per edge, output rows + rowptr per node
    nbytes = E * (4 * F * (ca if A is not None else 0) + (4 * F if W is not None else 0)
                  + (4 * C if U is not None else 0) + (4 if A is not None else 0)
                  + (4 if perm is not None else 0)) + N * (4 * CO * F + 4)
    # compulsory HBM bytes: the gathered table crosses HBM once (it is re-read from L2 afterwards)
    table = 0 if A is None else A.size(0) * 4 * F * ca
    dram = nbytes - E * 4 * F * (ca if A is not None else 0) + table
    _hgk.call("hgk_contract_node", _hgk.ptr(out), CO * F, CO, _hgk.ptr(a_t), lda, ca, _hgk.ptr(w_t), ldw,
              _hgk.ptr(u_t), _hgk.ptr(rowptr), _hgk.ptr(perm), _hgk.ptr(gidx), N, E, F, C,
              cost=(nbytes, 0, dram))
    return out


def _edge_operands(topo, G, side_g, X, side_x):
    g_t, ldg, cg = None, 0, 1
    if G is not None:
        g_t, ldg = _hgk.rows(G)
        cg = _channels(G)
    x_t, ldx, cx = None, 0, 1
    if X is not None:
        x_t, ldx = _hgk.rows(X)
        cx = _channels(X)
    gi = topo.idx[side_g] if G is not None else None
    xi = topo.idx[side_x] if X is not None else None
    return g_t, ldg, cg, gi, x_t, ldx, cx, xi


def _edge_w_raw(topo, G, side_g, X, side_x, U, C, F):
    out = torch.empty((topo.E, F), dtype=torch.float32, device=topo.device)
    g_t, ldg, cg, gi, x_t, ldx, cx, xi = _edge_operands(topo, G, side_g, X, side_x)
    u_t = None if U is None else U.contiguous()
    E = topo.E
    nbytes = E * (4 * F * ((cg if G is not None else 0) + (cx if X is not None else 0)) + 4 * F
                  + (4 * C if U is not None else 0) + 4 * ((G is not None) + (X is not None)))
    gathered = E * 4 * F * ((cg if G is not None else 0) + (cx if X is not None else 0))
    tables = (0 if G is None else G.size(0) * 4 * F * cg) + (0 if X is None else X.size(0) * 4 * F * cx)
    _hgk.call("hgk_contract_edge_w", _hgk.ptr(out), F, _hgk.ptr(g_t), ldg, cg, _hgk.ptr(gi),
              _hgk.ptr(x_t), ldx, cx, _hgk.ptr(xi), _hgk.ptr(u_t), E, F, C,
              cost=(nbytes, 0, nbytes - gathered + tables))
    return out


def _edge_u_raw(topo, G, side_g, X, side_x, W, C, F):
    out = torch.empty((topo.E, C), dtype=torch.float32, device=topo.device)
    g_t, ldg, cg, gi, x_t, ldx, cx, xi = _edge_operands(topo, G, side_g, X, side_x)
    w_t, ldw = (None, 0) if W is None else _hgk.rows(W)
    E = topo.E
    nbytes = E * (4 * F * ((cg if G is not None else 0) + (cx if X is not None else 0))
                  + (4 * F if W is not None else 0) + 4 * C + 4 * ((G is not None) + (X is not None)))
    _hgk.call("hgk_contract_edge_u", _hgk.ptr(out), _hgk.ptr(g_t), ldg, cg, _hgk.ptr(gi),
              _hgk.ptr(x_t), ldx, cx, _hgk.ptr(xi), _hgk.ptr(w_t), ldw, E, F, C, cost=(nbytes, 0))
    return out


class ContractNode(torch.autograd.Function):
    """out[n,c,f] = sum_{e: idx[side_out][e]=n} A[idx[side_a][e],c|0,f] W[e,f] U[e,c]."""

    @staticmethod
    def forward(ctx, A, W, U, topo, side_out, side_a, C, CO, F):
        ctx.topo, ctx.side_out, ctx.side_a, ctx.C, ctx.CO, ctx.F = topo, side_out, side_a, C, CO, F
        ctx.save_for_backward(A, W, U)
        return _node_raw(topo, side_out, A, side_a, W, U, C, CO, F)

    @staticmethod
    def backward(ctx, g):
        A, W, U = ctx.saved_tensors
        topo, so, sa, C, F = ctx.topo, ctx.side_out, ctx.side_a, ctx.C, ctx.F
        gA = gW = gU = None
        if A is not None and ctx.needs_input_grad[0]:
            gA = ContractNode.apply(g, W, U, topo, sa, so, C, _channels(A), F)
        if W is not None and ctx.needs_input_grad[1]:
            gW = ContractEdgeW.apply(g, A, U, topo, so, sa, C, F)
        if U is not None and ctx.needs_input_grad[2]:
            gU = ContractEdgeU.apply(g, A, W, topo, so, sa, C, F)
        return gA, gW, gU, None, None, None, None, None, None


class ContractEdgeW(torch.autograd.Function):
    """EW[e,f] = sum_c G[idx[side_g][e],c|0,f] X[idx[side_x][e],c|0,f] U[e,c]."""

    @staticmethod
    def forward(ctx, G, X, U, topo, side_g, side_x, C, F):
        ctx.topo, ctx.side_g, ctx.side_x, ctx.C, ctx.F = topo, side_g, side_x, C, F
        ctx.save_for_backward(G, X, U)
        return _edge_w_raw(topo, G, side_g, X, side_x, U, C, F)

    @staticmethod
    def backward(ctx, g):
        G, X, U = ctx.saved_tensors
        topo, sg, sx, C, F = ctx.topo, ctx.side_g, ctx.side_x, ctx.C, ctx.F
        gG = gX = gU = None
        if G is not None and ctx.needs_input_grad[0]:
            gG = ContractNode.apply(X, g, U, topo, sg, sx, C, _channels(G), F)
        if X is not None and ctx.needs_input_grad[1]:
            gX = ContractNode.apply(G, g, U, topo, sx, sg, C, _channels(X), F)
        if U is not None and ctx.needs_input_grad[2]:
            gU = ContractEdgeU.apply(G, X, g, topo, sg, sx, C, F)
        return gG, gX, gU, None, None, None, None, None


class ContractEdgeU(torch.autograd.Function):
    """EU[e,c] = sum_f G[idx[side_g][e],c|0,f] X[idx[side_x][e],c|0,f] W[e,f]."""

    @staticmethod
    def forward(ctx, G, X, W, topo, side_g, side_x, C, F):
        ctx.topo, ctx.side_g, ctx.side_x, ctx.C, ctx.F = topo, side_g, side_x, C, F
        ctx.save_for_backward(G, X, W)
        return _edge_u_raw(topo, G, side_g, X, side_x, W, C, F)

    @staticmethod
    def backward(ctx, g):
        G, X, W = ctx.saved_tensors
        topo, sg, sx, C, F = ctx.topo, ctx.side_g, ctx.side_x, ctx.C, ctx.F
        gG = gX = gW = None
        if G is not None and ctx.needs_input_grad[0]:
            gG = ContractNode.apply(X, W, g, topo, sg, sx, C, _channels(G), F)
        if X is not None and ctx.needs_input_grad[1]:
            gX = ContractNode.apply(G, W, g, topo, sx, sg, C, _channels(X), F)
        if W is not None and ctx.needs_input_grad[2]:
            gW = ContractEdgeW.apply(G, X, g, topo, sg, sx, C, F)
        return gG, gX, gW, None, None, None, None, None


# ---------------------------------------------------------------- functional --
def _feat(*ts):
    for t in ts:
        if t is not None:
            return t.size(-1)
    raise ValueError("cannot infer feature width: pass F")


def contract_node(topo, side_out, A=None, side_a=None, W=None, U=None, C=1, CO=None, F=None):
    if side_a is None:
        side_a = 1 - side_out
    if F is None:
        F = _feat(A, W)
    if CO is None:
        CO = C
    return ContractNode.apply(A, W, U, topo, side_out, side_a, C, CO, F)


def contract_edge_w(topo, G=None, side_g=0, X=None, side_x=1, U=None, C=1, F=None):
    if F is None:
        F = _feat(G, X)
    return ContractEdgeW.apply(G, X, U, topo, side_g, side_x, C, F)


def contract_edge_u(topo, G=None, side_g=0, X=None, side_x=1, W=None, C=1, F=None):
    if F is None:
        F = _feat(G, X, W)
    return ContractEdgeU.apply(G, X, W, topo, side_g, side_x, C, F)


def segment_sum(edge_values, topo, side):
    """out[n,:] = sum of edge_values[e,:] over edges with idx[side][e] == n.

    Replaces unsorted_segment_sum (EGCLStack.py:292-298), index_add_
    (PAINNStack.py:256) and scatter_add (Base.py:598)."""
    return contract_node(topo, side, W=edge_values, F=edge_values.size(-1))


def segment_mean(edge_values, topo, side):
    """unsorted_segment_mean (utils/model/model.py:279-286): sum / clamp(count, 1)."""
    total = segment_sum(edge_values, topo, side)
    count = topo.degree(side).clamp(min=1).to(total.dtype).unsqueeze(-1)
    return total / count


def gather(node_values, topo, side):
    """edge[e,:] = node_values[idx[side][e], :] with a deterministic backward."""
    return contract_edge_w(topo, G=node_values, side_g=side, F=node_values.size(-1))


def gather_mul_segment_sum(x, W, topo, side_out=1, side_in=0):
    """out[t] = sum_{e: idx[side_out][e]=t} x[idx[side_in][e]] * W[e]  (CFConv, SCFStack.py:285-293)."""
    return contract_node(topo, side_out, A=x, side_a=side_in, W=W, F=W.size(-1))


def _edge_gather_add_raw(topo, P, side_p, Q, side_q):
    F = P.size(-1)
    out = torch.empty((topo.E, F), dtype=torch.float32, device=P.device)
    p_t, ldp = _hgk.rows(P)
    q_t, ldq = _hgk.rows(Q)
    _hgk.call("hgk_edge_gather_add", _hgk.ptr(out), F, _hgk.ptr(p_t), ldp, _hgk.ptr(topo.idx[side_p]),
              _hgk.ptr(q_t), ldq, _hgk.ptr(topo.idx[side_q]), topo.E, F,
              cost=(topo.E * (3 * 4 * F + 8), 0))
    return out


class EdgeGatherAdd(torch.autograd.Function):
    """out[e,:] = P[idx[side_p][e],:] + Q[idx[side_q][e],:]  (split-weight edge-MLP input)."""

    @staticmethod
    def forward(ctx, P, Q, topo, side_p, side_q):
        ctx.topo, ctx.side_p, ctx.side_q = topo, side_p, side_q
        return _edge_gather_add_raw(topo, P, side_p, Q, side_q)

    @staticmethod
    def backward(ctx, g):
        gP = segment_sum(g, ctx.topo, ctx.side_p) if ctx.needs_input_grad[0] else None
        gQ = segment_sum(g, ctx.topo, ctx.side_q) if ctx.needs_input_grad[1] else None
        return gP, gQ, None, None, None


def edge_gather_add(P, Q, topo, side_p=0, side_q=1):
    return EdgeGatherAdd.apply(P, Q, topo, side_p, side_q)
