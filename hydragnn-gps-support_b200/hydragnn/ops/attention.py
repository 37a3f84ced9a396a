"""Variable-length multi-head attention core (hgk_attn_varlen_fwd / _bwd).

Replaces the core of the reference's GPS global attention
(hydragnn/globalAtt/gps.py:126-133): to_dense_batch -> nn.MultiheadAttention ->
boolean unpad.  q, k, v are the already projected [N, H*D] node matrices; the
segments (`cu_seqlens`) say which nodes attend to each other.
"""
import math

import torch

from .. import _hgk


def _sum_sq(cu_seqlens):
    """sum of squared segment lengths (host value cached on the tensor: no sync in the hot loop)."""
    v = getattr(cu_seqlens, "_hgk_sum_sq", None)
    if v is None:
        lens = (cu_seqlens[1:] - cu_seqlens[:-1]).to(torch.int64)
        v = int((lens * lens).sum())
        try:
            cu_seqlens._hgk_sum_sq = v
        except Exception:
            pass
    return v


def _attn_fwd_raw(q, k, v, cu_seqlens, H, scale):
    N, HD = q.shape
    D = HD // H
    q_t, ldq = _hgk.rows(q)
    k_t, ldk = _hgk.rows(k)
    v_t, ldv = _hgk.rows(v)
    o = torch.empty((N, HD), dtype=torch.float32, device=q.device)
    lse = torch.empty((N, H), dtype=torch.float32, device=q.device)
    S = cu_seqlens.numel() - 1
    _hgk.call("hgk_attn_varlen_fwd", _hgk.ptr(o), HD, _hgk.ptr(lse), _hgk.ptr(q_t), ldq, _hgk.ptr(k_t), ldk,
              _hgk.ptr(v_t), ldv, _hgk.ptr(cu_seqlens), S, N, H, D, float(scale),
              cost=(4 * (4 * N * HD + N * H), 4 * HD * _sum_sq(cu_seqlens)))
    return o, lse, q_t, k_t, v_t


def _attn_bwd_raw(do, o, lse, q, k, v, cu, H, scale):
    N, HD = o.shape
    D = HD // H
    do_t, lddo = _hgk.rows(do)
    dq = torch.empty((N, HD), dtype=torch.float32, device=o.device)
    dk = torch.empty_like(dq)
    dv = torch.empty_like(dq)
    delta = torch.empty((N, H), dtype=torch.float32, device=o.device)
    q_t, ldq = _hgk.rows(q)
    k_t, ldk = _hgk.rows(k)
    v_t, ldv = _hgk.rows(v)
    _hgk.call("hgk_attn_varlen_bwd", _hgk.ptr(dq), _hgk.ptr(dk), _hgk.ptr(dv), _hgk.ptr(delta),
              _hgk.ptr(do_t), lddo, _hgk.ptr(o), HD, _hgk.ptr(lse), _hgk.ptr(q_t), ldq, _hgk.ptr(k_t), ldk,
              _hgk.ptr(v_t), ldv, _hgk.ptr(cu), cu.numel() - 1, N, H, D, float(scale),
              cost=(4 * (8 * N * HD + 2 * N * H), 10 * HD * _sum_sq(cu)))
    return dq, dk, dv


def _dense_backward(do, q, k, v, cu, H, scale):
    """(dq, dk, dv) from differentiable torch ops, segment by segment.  Only used when the backward
    pass is itself being differentiated (create_graph=True: a force loss on a model whose attention
    inputs depend on the positions) -- the kernels provide first derivatives only.  Reads the
    segment boundaries on the host (one sync) and materialises the [H, L, L] score block per segment."""
    N, HD = q.shape
    D = HD // H
    bounds = [int(c) for c in cu.tolist()]
    dq, dk, dv = [], [], []
    for b, e in zip(bounds[:-1], bounds[1:]):
        L = e - b
        qs, ks, vs, gs = (t[b:e].reshape(L, H, D).transpose(0, 1) for t in (q, k, v, do))      # [H, L, D]
        p = torch.softmax((qs @ ks.transpose(1, 2)) * scale, dim=-1)
        dp = gs @ vs.transpose(1, 2)
        ds = p * (dp - (dp * p).sum(dim=-1, keepdim=True))
        dq.append(((ds @ ks) * scale).transpose(0, 1).reshape(L, HD))
        dk.append(((ds.transpose(1, 2) @ qs) * scale).transpose(0, 1).reshape(L, HD))
        dv.append((p.transpose(1, 2) @ gs).transpose(0, 1).reshape(L, HD))
    if not dq:
        z = q * 0.0
        return z, z, z
    return torch.cat(dq), torch.cat(dk), torch.cat(dv)


class AttnVarlen(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, cu_seqlens, H, scale):
        o, lse, q_t, k_t, v_t = _attn_fwd_raw(q, k, v, cu_seqlens, H, scale)
        ctx.save_for_backward(q, k, v, o, lse, cu_seqlens)
        ctx.H, ctx.scale = H, scale
        return o

    @staticmethod
    def backward(ctx, do):
        q, k, v, o, lse, cu = ctx.saved_tensors
        if torch.is_grad_enabled():        # create_graph=True: stay differentiable
            dq, dk, dv = _dense_backward(do, q, k, v, cu, ctx.H, ctx.scale)
        else:
            dq, dk, dv = _attn_bwd_raw(do, o, lse, q, k, v, cu, ctx.H, ctx.scale)
        return dq, dk, dv, None, None, None


def attn_varlen(q, k, v, cu_seqlens, num_heads, scale=None):
    """softmax(q k^T * scale) v inside each segment; q,k,v [N, H*D] fp32."""
    D = q.size(1) // num_heads
    if scale is None:
        scale = 1.0 / math.sqrt(D)
    return AttnVarlen.apply(q, k, v, cu_seqlens.to(torch.int32), num_heads, scale)
