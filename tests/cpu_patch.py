"""Test-only: run the product's host logic on CPU by swapping the C-ABI launchers
(hydragnn.ops.*._xxx_raw) for their CPU restatement in oracle/kernel_spec.py.

The product itself has no CPU path (hydragnn._hgk refuses CPU tensors).  This patch exists so
that `-m "not gpu"` tests can check everything ABOVE the kernels -- weight splitting, which
CSR side each conv aggregates on, the autograd closure used by the force loss -- against the
golden reference outputs in a container without a GPU.
"""
import contextlib

import torch

from oracle import kernel_spec as ks


def _node_raw(topo, side_out, A, side_a, W, U, C, CO, F):
    gidx = topo.idx[side_a] if A is not None else None
    return ks.contract_node(topo.idx[side_out], topo.num_rows[side_out], A, gidx, W, U, C, CO, F)


def _edge_w_raw(topo, G, side_g, X, side_x, U, C, F):
    return ks.contract_edge_w(G, topo.idx[side_g] if G is not None else None,
                              X, topo.idx[side_x] if X is not None else None, U, topo.E, C, F)


def _edge_u_raw(topo, G, side_g, X, side_x, W, C, F):
    return ks.contract_edge_u(G, topo.idx[side_g] if G is not None else None,
                              X, topo.idx[side_x] if X is not None else None, W, topo.E, C, F)


def _edge_gather_add_raw(topo, P, side_p, Q, side_q):
    return ks.edge_gather_add(P, topo.idx[side_p], Q, topo.idx[side_q])


def _pna_operands(P, Q, R, topo, side):
    """kernel_spec's functions take dense P / Q and aggregate at `dst`: zeros for absent operands,
    endpoints swapped for side 0."""
    first = next(t for t in (P, Q, R) if t is not None)
    N, F = topo.num_rows[side], first.size(1)
    zero = first.new_zeros((N, F))
    return (zero if P is None else P, zero if Q is None else Q, topo.idx[1 - side], topo.idx[side], N)


def _pna_fwd_raw(P, Q, R, topo, M=None, side=1):
    p, q, src, dst, N = _pna_operands(P, Q, R, topo, side)
    return ks.pna_aggregate_fwd(p, q, R, src, dst, N, M)


def _pna_bwd_raw(gout, out, cmin, cmax, P, Q, R, topo, M=None, side=1):
    p, q, src, dst, N = _pna_operands(P, Q, R, topo, side)
    res = ks.pna_aggregate_bwd(gout, out, cmin, cmax, p, q, R, src, dst, N, M)
    return res if M is not None else (res, None)


def _attn_fwd_raw(q, k, v, cu, H, scale):
    o, lse = ks.attn_varlen_fwd(q, k, v, cu, H, scale)
    return o, lse, q, k, v


def _attn_bwd_raw(do, o, lse, q, k, v, cu, H, scale, packed=False):
    grads = ks.attn_varlen_bwd(do, q, k, v, cu, H, scale)
    return torch.cat(grads, dim=1) if packed else grads


def _build_csr(key, num_rows):
    return ks.csr_build(key, num_rows)


def _radius_graph(pos, ptr, r, max_num_neighbors=0, loop=False, cell=None, pbc=None, mode=None,
                  return_rowptr=False):
    if mode is None:
        mode = 0 if cell is not None else 1
    src, dst, S, dist, rowptr = ks.radius_graph(pos, ptr, r, int(max_num_neighbors or 0), loop,
                                                cell, pbc, mode)
    out = {"edge_index": torch.stack([src, dst]).to(torch.int64), "src": src, "dst": dst,
           "shift_int": S, "dist": dist, "rowptr": rowptr}
    if cell is not None:
        g = torch.bucketize(dst.long(), torch.as_tensor(ptr)[1:].long(), right=True)
        out["edge_shifts"] = torch.einsum("ek,ekd->ed", S.float(), cell.float().reshape(-1, 3, 3)[g])
    return out


# ---- fused SchNet filter network (hydragnn/ops/filter.py) ----
def _filter_fwd_raw(dist, raw, offset, coeff, cutoff, w0p, b0, w1, b1):
    return ks.filter_fwd_padded(dist, raw, offset, coeff, cutoff, w0p, b0, w1, b1)


def _filter_bwd_raw(gw, h, dist, cutoff, w1t):
    return ks.filter_bwd(gw, h, dist, cutoff, w1t)


def _wgrad_raw(a, scale_d, cutoff, b=None, gauss=None):
    return ks.wgrad_edges(a, scale_d, cutoff, b, gauss)


def _cfconv_filter_bwd_raw(g, x1, topo, h, dist, cutoff, w1t):
    gw = g[topo.idx[1].long()] * x1[topo.idx[0].long()]
    return ks.filter_bwd(gw, h, dist, cutoff, w1t)


def _cfconv_wgrad_raw(g, x1, topo, dist, cutoff, h):
    gw = g[topo.idx[1].long()] * x1[topo.idx[0].long()]
    return ks.wgrad_edges(gw, dist, cutoff, b=h)


def _fused_filter_ok(dist, raw, num_filters, num_gaussians):
    from hydragnn.ops import filter as filt
    if dist.requires_grad or dist.dtype != torch.float32 or (raw is not None and raw.requires_grad):
        return False
    r = 0 if raw is None else raw.size(1)
    return (num_filters == filt.SUPPORTED["F"] and num_gaussians == filt.SUPPORTED["G"] and r in filt.SUPPORTED["R"])


@contextlib.contextmanager
def cpu_kernels():
    from hydragnn.ops import attention, contract, geometry, node, pna, topo
    from hydragnn.ops import filter as filt
    import importlib

    import hydragnn.ops as ops_pkg

    scf_mod = importlib.import_module("hydragnn.models.SCFStack")
    radius_mod = importlib.import_module("hydragnn.ops.radius")

    patches = [
        (contract, "_node_raw", _node_raw), (contract, "_edge_w_raw", _edge_w_raw),
        (contract, "_edge_u_raw", _edge_u_raw), (contract, "_edge_gather_add_raw", _edge_gather_add_raw),
        (pna, "_pna_fwd_raw", _pna_fwd_raw), (pna, "_pna_bwd_raw", _pna_bwd_raw),
        (attention, "_attn_fwd_raw", _attn_fwd_raw), (attention, "_attn_bwd_raw", _attn_bwd_raw),
        (topo, "build_csr", _build_csr), (ops_pkg, "radius_graph", _radius_graph),
        # node-row kernels of the GPS block: the autograd Functions of ops/node.py run over their restatement
        (node, "linear_tc", ks.linear_tc), (node, "wgrad_rows", ks.wgrad_rows), (node, "bn_stats", ks.bn_stats),
        (node, "bn_apply", ks.bn_apply), (node, "bn_bwd", ks.bn_bwd), (node, "mask_axpy", ks.mask_axpy),
        (node, "device_ok", lambda t: t.dtype == torch.float32),
        (geometry, "_fwd_raw", ks.edge_geom_fwd), (geometry, "_bwd_raw", ks.edge_geom_bwd),
        (geometry, "_bwd2_raw", ks.edge_geom_bwd2),
        (filt, "_filter_fwd_raw", _filter_fwd_raw), (filt, "_filter_bwd_raw", _filter_bwd_raw),
        (filt, "_wgrad_raw", _wgrad_raw), (filt, "_cfconv_filter_bwd_raw", _cfconv_filter_bwd_raw),
        (filt, "_cfconv_wgrad_raw", _cfconv_wgrad_raw), (filt, "fused_filter_ok", _fused_filter_ok),
        (scf_mod, "fused_filter_ok", _fused_filter_ok),
        (scf_mod, "radius_graph", _radius_graph), (radius_mod, "radius_graph", _radius_graph),
    ]
    saved = [(m, n, getattr(m, n)) for m, n, _ in patches]
    try:
        for m, n, f in patches:
            setattr(m, n, f)
        yield
    finally:
        for m, n, f in saved:
            setattr(m, n, f)
