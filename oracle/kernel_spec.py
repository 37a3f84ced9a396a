"""ORACLE (test infrastructure only -- never imported by the product path).

CPU restatement of every entry point of include/hgk.h in terms of the torch CPU ops the
reference issues at the call site that entry point replaces (index_select, sequential
index_add_/scatter_add_, dense softmax attention, scatter_reduce).  Two uses, both in
tests/:
  * `-m gpu` tests call the C ABI and compare with these functions on the same inputs
    (bit-exact where hgk.h promises "segment order", a stated tolerance elsewhere);
  * `-m "not gpu"` tests patch these functions under the product's autograd wrappers
    (tests/cpu_patch.py) so the host-side logic of the stacks -- split weights, CSR sides,
    double backward -- is checked against the golden reference outputs without a GPU.
"""
import math

import numpy as np
import torch

from . import pyg_lite as pyg
from . import hydragnn_ref as ref


# hgk_csr_build: stable grouping of edge ids by key (a sequential scatter visits e = 0..E-1)
def csr_build(key, num_rows):
    key = key.long()
    counts = torch.bincount(key, minlength=num_rows)
    rowptr = torch.zeros(num_rows + 1, dtype=torch.int32)
    rowptr[1:] = torch.cumsum(counts, 0)
    perm = torch.sort(key, stable=True).indices.to(torch.int32)
    return rowptr, perm


def _bc(t, C):
    """[rows, F] or [rows, CA, F] -> [rows, C, F] view."""
    if t.dim() == 2:
        t = t.unsqueeze(1)
    return t.expand(t.size(0), C, t.size(2))


# hgk_contract_node -- index_select + mul (+ mul) + index_add_  (SCFStack.py:285-293,
# EGCLStack.py:292-298, PAINNStack.py:238-257, Base.py:478,598)
def contract_node(key, num_rows, A, gidx, W, U, C, CO, F):
    E = key.numel()
    term = None
    if A is not None:
        term = _bc(A.index_select(0, gidx.long()), C)
    if W is not None:
        w = W.reshape(E, 1, F).expand(E, C, F)
        term = w if term is None else term * w
    if U is not None:
        u = U.reshape(E, C, 1)
        term = u.expand(E, C, F) if term is None else term * u
    term = term.contiguous()
    if CO == C:
        out = torch.zeros((num_rows, C, F), dtype=torch.float32)
        out.index_add_(0, key.long(), term)
        return out.reshape(num_rows, F) if C == 1 else out
    out = torch.zeros((num_rows, F), dtype=torch.float32)          # CO == 1 < C: also sum over c
    out.index_add_(0, key.long().repeat_interleave(C), term.reshape(E * C, F))
    return out


# hgk_contract_edge_w -- per-edge product of two gathered rows (backward of the above w.r.t. W)
def contract_edge_w(G, gi, X, xi, U, E, C, F):
    term = None
    if G is not None:
        term = _bc(G.index_select(0, gi.long()), C)
    if X is not None:
        x = _bc(X.index_select(0, xi.long()), C)
        term = x if term is None else term * x
    if U is not None:
        term = term * U.reshape(E, C, 1)
    acc = term[:, 0]
    for c in range(1, C):
        acc = acc + term[:, c]
    return acc.contiguous()


# hgk_contract_edge_u -- per-edge dot product over features (fixed tree on the GPU; plain sum here)
def contract_edge_u(G, gi, X, xi, W, E, C, F):
    term = None
    if G is not None:
        term = _bc(G.index_select(0, gi.long()), C)
    if X is not None:
        x = _bc(X.index_select(0, xi.long()), C)
        term = x if term is None else term * x
    if W is not None:
        term = term * W.reshape(E, 1, F)
    return term.sum(dim=-1).contiguous()


# hgk_edge_gather_add -- x[row] W_a + x[col] W_b of E_GCL.edge_model in split-weight form
def edge_gather_add(P, pi, Q, qi):
    out = P.index_select(0, pi.long())
    if Q is not None:
        out = out + Q.index_select(0, qi.long())
    return out


# hgk_pna_aggregate_fwd -- PyG PNAConv.propagate + aggregators mean,min,max,std (PNAStack.py:30-53)
def pna_aggregate_fwd(P, Q, R, src, dst, N, M=None):
    F = P.size(1)
    h = P.index_select(0, dst.long()) + Q.index_select(0, src.long())
    if R is not None:
        h = h + R
    if M is not None:            # PNAPlus: Hadamard product with the projected radial basis
        h = h * M
    idx = dst.long()
    mean = pyg.scatter(h, idx, 0, N, "mean")
    mn = pyg.scatter(h, idx, 0, N, "min")
    mx = pyg.scatter(h, idx, 0, N, "max")
    mean2 = pyg.scatter(h * h, idx, 0, N, "mean")
    sd = (mean2 - mean * mean).clamp(min=1e-5).sqrt()
    sd = sd.masked_fill(sd <= math.sqrt(1e-5), 0.0)
    deg = torch.bincount(idx, minlength=N)
    sd = sd * (deg > 0).unsqueeze(1)
    cmin = torch.zeros((N, F), dtype=torch.int32).index_add_(0, idx, (h == mn[idx]).to(torch.int32))
    cmax = torch.zeros((N, F), dtype=torch.int32).index_add_(0, idx, (h == mx[idx]).to(torch.int32))
    return torch.cat([mean, mn, mx, sd], dim=1), cmin, cmax


# hgk_pna_aggregate_bwd -- d/dh of the four aggregators (ties split evenly like scatter_reduce)
def pna_aggregate_bwd(gout, out, cmin, cmax, P, Q, R, src, dst, N, M=None):
    """-> gh, or (gh, gm) when the multiplier M is given."""
    F = P.size(1)
    idx = dst.long()
    h = P.index_select(0, idx) + Q.index_select(0, src.long())
    if R is not None:
        h = h + R
    pre = h
    if M is not None:
        h = h * M
    cnt = torch.bincount(idx, minlength=N).to(torch.float32).unsqueeze(1)[idx]
    o, g = out[idx], gout[idx]
    gh = g[:, :F] / cnt
    gh = gh + torch.where(h == o[:, F:2 * F], g[:, F:2 * F] / cmin[idx].float(), torch.zeros_like(h))
    gh = gh + torch.where(h == o[:, 2 * F:3 * F], g[:, 2 * F:3 * F] / cmax[idx].float(), torch.zeros_like(h))
    sd = o[:, 3 * F:]
    safe = torch.where(sd > 0, sd, torch.ones_like(sd))
    gh = gh + torch.where(sd > 0, g[:, 3 * F:] * (h - o[:, :F]) / (cnt * safe), torch.zeros_like(h))
    if M is not None:
        return gh * M, gh * pre
    return gh


# hgk_attn_varlen_fwd/_bwd -- to_dense_batch + MultiheadAttention core + unpad (gps.py:126-133)
def attn_varlen_fwd(q, k, v, cu_seqlens, H, scale):
    N, HD = q.shape
    D = HD // H
    o = torch.zeros((N, HD), dtype=q.dtype)        # fp32 restatement; fp64 inputs give the tests' truth
    lse = torch.zeros((N, H), dtype=q.dtype)
    cu = [int(c) for c in cu_seqlens]
    for b, e in zip(cu[:-1], cu[1:]):
        for h in range(H):
            sl = slice(h * D, (h + 1) * D)
            s = (q[b:e, sl] @ k[b:e, sl].t()) * scale
            lse[b:e, h] = torch.logsumexp(s, dim=1)
            o[b:e, sl] = torch.softmax(s, dim=1) @ v[b:e, sl]
    return o, lse


def attn_varlen_bwd(do, q, k, v, cu_seqlens, H, scale):
    q, k, v = (t.detach().clone().requires_grad_(True) for t in (q, k, v))
    with torch.enable_grad():
        o, _ = attn_varlen_fwd_autograd(q, k, v, cu_seqlens, H, scale)
        dq, dk, dv = torch.autograd.grad(o, (q, k, v), do)
    return dq, dk, dv


def attn_varlen_fwd_autograd(q, k, v, cu_seqlens, H, scale):
    N, HD = q.shape
    D = HD // H
    cu = [int(c) for c in cu_seqlens]
    rows = []
    for b, e in zip(cu[:-1], cu[1:]):
        heads = []
        for h in range(H):
            sl = slice(h * D, (h + 1) * D)
            s = (q[b:e, sl] @ k[b:e, sl].t()) * scale
            heads.append(torch.softmax(s, dim=1) @ v[b:e, sl])
        rows.append(torch.cat(heads, dim=1))
    return (torch.cat(rows, dim=0) if rows else q.new_zeros((0, HD))), None


# hgk_radius_graph_{count,fill}: the two neighbour searches of the reference, per graph
def radius_graph(pos, ptr, r, max_nbr, loop, cell=None, pbc=None, mode=1):
    """-> src, dst (int32, global node ids), shift_int [E,3], dist fp64 [E], rowptr int32 [N+1].
    mode 0 = ASE semantics (RadiusGraphPBC, graph_samples_checks_and_updates.py:146-282),
    mode 1 = torch_cluster semantics (RadiusGraph / RadiusInteractionGraph)."""
    pos = pos.detach().to(torch.float32).cpu()
    ptr = [int(p) for p in ptr]
    srcs, dsts, shifts, dists = [], [], [], []
    for g, (b, e) in enumerate(zip(ptr[:-1], ptr[1:])):
        p = pos[b:e]
        if mode == 1:
            ei = pyg.radius_graph_cluster(p, r, None, bool(loop), max_nbr if max_nbr and max_nbr > 0 else 0)
            x = p.numpy().astype(np.float32)
            d = x[ei[0].numpy()] - x[ei[1].numpy()]
            sq = (d * d).astype(np.float32)
            d2 = ((sq[:, 0] + sq[:, 1]).astype(np.float32) + sq[:, 2]).astype(np.float32)
            S = np.zeros((ei.size(1), 3), dtype=np.int64)
            dd = d2.astype(np.float64)
            ei = ei.numpy()
        else:
            if cell is None:
                c, pb = np.eye(3), np.zeros(3, dtype=bool)
            else:
                c, pb = cell[g].numpy(), np.asarray(pbc[g]).astype(bool)
            ei, S, dd = ref.radius_graph_pbc(p.numpy(), c, pb, r, max_nbr, bool(loop))
        srcs.append(ei[0] + b); dsts.append(ei[1] + b); shifts.append(S); dists.append(dd)
    src = np.concatenate(srcs) if srcs else np.zeros(0, np.int64)
    dst = np.concatenate(dsts) if dsts else np.zeros(0, np.int64)
    N = pos.size(0)
    rowptr = np.zeros(N + 1, dtype=np.int32)
    np.cumsum(np.bincount(dst, minlength=N), out=rowptr[1:])
    return (torch.from_numpy(src.astype(np.int32)), torch.from_numpy(dst.astype(np.int32)),
            torch.from_numpy(np.concatenate(shifts).astype(np.int32).reshape(-1, 3)),
            torch.from_numpy(np.concatenate(dists).astype(np.float64)), torch.from_numpy(rowptr))


# hgk_schnet_filter_fwd -- GaussianSmearing -> cat -> Linear -> ShiftedSoftplus -> Linear -> cosine
# envelope, exactly the op sequence of CFConv.forward (SCFStack.py:259-272) on the unfolded inputs
def schnet_filter_fwd(dist, raw, offset, coeff, cutoff, w0g, w0r, b0, w1, b1):
    delta = dist.view(-1, 1) - offset.view(1, -1)
    feats = torch.exp(coeff * torch.pow(delta, 2))
    weight = w0g
    if raw is not None:
        feats = torch.cat([feats, raw], dim=-1)
        weight = torch.cat([w0g, w0r], dim=1)
    h = torch.nn.functional.softplus(torch.nn.functional.linear(feats, weight, b0)) - math.log(2.0)
    env = 0.5 * (torch.cos(dist * math.pi / cutoff) + 1.0)
    return torch.nn.functional.linear(h, w1, b1) * env.view(-1, 1), h


# ---- node-row kernels of the GPS block (csrc/node_tc.cu; reference hydragnn/globalAtt/gps.py:72-83,116-150) ----
def linear_tc(x, w, bias=None, trans_w=False, relu=False, mask=None, nz=None, scale=1.0, residual=None):
    """hgk_linear_tc: residual + scale * mask * [nz != 0] * act(x w^T + bias) as the torch ops it replaces."""
    y = x @ (w if trans_w else w.t())
    if bias is not None:
        y = y + bias
    if relu:
        y = torch.relu(y)
    if mask is not None:
        y = y * mask.to(y.dtype) * scale
    if nz is not None:
        y = y * (nz != 0).to(y.dtype) * scale
    if residual is not None:
        y = y + residual
    return y


def wgrad_rows(a, b, want_bias=True):
    """hgk_wgrad_rows: (a^T b, column sums of a)."""
    return a.t() @ b, (a.sum(0) if want_bias else None)


def bn_stats(t, running_mean=None, running_var=None, num_batches_tracked=None, momentum=0.1, eps=1e-5, n_valid=None):
    """hgk_bn_stats: what nn.BatchNorm1d computes in training (biased variance normalises, the unbiased
    one goes into running_var); with n_valid, over the first n_valid rows only."""
    if n_valid is not None:
        t = t[:int(n_valid)]
    n = t.size(0)
    mean = t.mean(0)
    var = t.var(0, unbiased=False)
    if running_mean is not None:
        running_mean.mul_(1 - momentum).add_(momentum * mean)
        running_var.mul_(1 - momentum).add_(momentum * var * (n / max(n - 1, 1)))
    if num_batches_tracked is not None:
        num_batches_tracked.add_(1)
    return mean, torch.rsqrt(var + eps)


def _bn_side(t, p):
    y = (t - p.mean) * p.rstd
    if p.gamma is not None:
        y = y * p.gamma
    if p.beta is not None:
        y = y + p.beta
    return y


def bn_apply(ta, pa, tb=None, pb=None, relu=False):
    y = _bn_side(ta, pa)
    if tb is not None:
        y = y + _bn_side(tb, pb)
    return torch.relu(y) if relu else y


def bn_bwd(g, ta, pa, tb=None, pb=None, mask_a=None, scale_a=1.0, mask_b=None, scale_b=1.0, n_valid=None):
    """hgk_bn_bwd: the standard BatchNorm backward with batch statistics,
    dt = gamma rstd (g - mean(g) - xhat mean(g xhat)); with n_valid the sums run over the first n_valid rows
    and the rows beyond get zero."""
    n = ta.size(0) if n_valid is None else int(n_valid)
    rows = (torch.arange(ta.size(0)) < n).to(ta.dtype).unsqueeze(1)
    out = {}
    for key, t, p, mask, scale in (("a", ta, pa, mask_a, scale_a), ("b", tb, pb, mask_b, scale_b)):
        if t is None:
            continue
        xhat = (t - p.mean) * p.rstd
        s0, sx = (g * rows).sum(0), (g * xhat * rows).sum(0)
        dt = p.rstd * (g - s0 / n - xhat * (sx / n)) * rows
        if p.gamma is not None:
            dt = dt * p.gamma
        out["dt" + key] = dt
        out["dz" + key] = dt if mask is None else dt * mask.to(dt.dtype) * scale
        out["dgamma_" + key] = sx if p.gamma is not None else None
        out["dbeta_" + key] = s0
    return out


def mask_axpy(z, mask=None, scale=1.0, residual=None):
    y = z if mask is None else z * mask.to(z.dtype) * scale
    return y if residual is None else y + residual


# ---- launcher-level restatements of hydragnn/ops/filter.py (padded-weight layout of the fused kernels) ----
def _envelope(dist, cutoff):
    return 0.5 * (torch.cos(dist * math.pi / cutoff) + 1.0)


def _filter_input_padded(dist, raw, offset, coeff):
    """[gauss(d) | pad to a multiple of 4 | raw | pad to 64]: the input layout of hgk_schnet_filter_fwd."""
    G = offset.numel()
    GP = (G + 3) // 4 * 4
    x = dist.new_zeros((dist.numel(), 64))
    x[:, :G] = torch.exp(coeff * torch.pow(dist.view(-1, 1) - offset.view(1, -1), 2))
    if raw is not None:
        x[:, GP:GP + raw.size(1)] = raw
    return x


def filter_fwd_padded(dist, raw, offset, coeff, cutoff, w0p, b0, w1, b1):
    """hgk_schnet_filter_fwd on the padded first-layer weight w0p [F, 64] -> (w, h)."""
    h = torch.nn.functional.softplus(_filter_input_padded(dist, raw, offset, coeff) @ w0p.t() + b0) - math.log(2.0)
    return (h @ w1.t() + b1) * _envelope(dist, cutoff).view(-1, 1), h


def filter_bwd(gw, h, dist, cutoff, w1t):
    """hgk_schnet_filter_bwd: dZ0 = ((gw * envelope) W1) * sigmoid(Z0), sigmoid(z0) = 1 - exp(-(h + ln 2))."""
    dz1 = gw * _envelope(dist, cutoff).view(-1, 1)
    return (dz1 @ w1t.t()) * (1.0 - torch.exp(-(h + math.log(2.0))))


def wgrad_edges(a, scale_d, cutoff, b=None, gauss=None):
    """hgk_wgrad_tn + hgk_wgrad_reduce: (A^T B, column sums of A), A optionally scaled by the envelope,
    B = b or the padded filter input."""
    if scale_d is not None:
        a = a * _envelope(scale_d, cutoff).view(-1, 1)
    if b is None:
        b = _filter_input_padded(*gauss)
    return a.t() @ b, a.sum(0)


# ---- edge geometry with explicit derivatives (csrc/edge_geom.cu; reference utils/model/operations.py:21-36) ----
def edge_geom_fwd(pos, src, dst, shifts, eps, want_unit):
    vec = pos[dst.long()] - pos[src.long()]
    if shifts is not None:
        vec = vec + shifts
    length = torch.linalg.norm(vec, dim=-1)
    unit = vec / (length + eps).unsqueeze(-1) if want_unit else None
    return vec, length, unit


def _inv_len(length):
    return torch.where(length > 0, 1.0 / length.clamp(min=1e-300 if length.dtype == torch.float64 else 1e-38),
                       torch.zeros_like(length))


def edge_geom_bwd(g_vec, g_len, g_unit, vec, length, eps):
    """t = d loss / d vec given the gradients of (vec, len, unit)."""
    inv = _inv_len(length).unsqueeze(-1)
    t = torch.zeros_like(vec) if g_vec is None else g_vec.clone()
    if g_len is not None:
        t = t + g_len.unsqueeze(-1) * vec * inv
    if g_unit is not None:
        d = 1.0 / (length + eps).unsqueeze(-1)
        t = t + g_unit * d - vec * (g_unit * vec).sum(-1, keepdim=True) * inv * d * d
    return t


def edge_geom_bwd2(h, g_len, g_unit, vec, length, eps):
    """(d g_len, d g_unit, d vec) for h = d loss / d t (the derivative of edge_geom_bwd)."""
    L = length.unsqueeze(-1)
    inv = _inv_len(length).unsqueeze(-1)
    d = 1.0 / (L + eps)
    c = inv * d * d
    hv = (h * vec).sum(-1, keepdim=True)
    d_glen = (hv * inv).squeeze(-1)
    d_gunit = h * d - c * hv * vec
    dv = torch.zeros_like(vec)
    if g_len is not None:
        gl = g_len.unsqueeze(-1)
        dv = dv + gl * inv * h - gl * hv * inv ** 3 * vec
    if g_unit is not None:
        hg = (h * g_unit).sum(-1, keepdim=True)
        gv = (g_unit * vec).sum(-1, keepdim=True)
        cp = -inv * inv * d * d - 2.0 * inv * d ** 3
        dv = dv - ((hg * c + hv * gv * cp * inv) * vec + gv * c * h + hv * c * g_unit)
    return d_glen, d_gunit, dv
