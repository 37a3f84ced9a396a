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
    _, ldq = _hgk.rows(q)
    _, ldk = _hgk.rows(k)
    _, ldv = _hgk.rows(v)
    _hgk.call("hgk_attn_varlen_bwd", _hgk.ptr(dq), _hgk.ptr(dk), _hgk.ptr(dv), _hgk.ptr(delta),
              _hgk.ptr(do_t), lddo, _hgk.ptr(o), HD, _hgk.ptr(lse), _hgk.ptr(q), ldq, _hgk.ptr(k), ldk,
              _hgk.ptr(v), ldv, _hgk.ptr(cu), cu.numel() - 1, N, H, D, float(scale),
              cost=(4 * (8 * N * HD + 2 * N * H), 10 * HD * _sum_sq(cu)))
    return dq, dk, dv


class AttnVarlen(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, cu_seqlens, H, scale):
        o, lse, q_t, k_t, v_t = _attn_fwd_raw(q, k, v, cu_seqlens, H, scale)
        ctx.save_for_backward(q_t, k_t, v_t, o, lse, cu_seqlens)
        ctx.H, ctx.scale = H, scale
        return o

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, do):
        q, k, v, o, lse, cu = ctx.saved_tensors
        dq, dk, dv = _attn_bwd_raw(do, o, lse, q, k, v, cu, ctx.H, ctx.scale)
        return dq, dk, dv, None, None, None


def attn_varlen(q, k, v, cu_seqlens, num_heads, scale=None):
    """softmax(q k^T * scale) v inside each segment; q,k,v [N, H*D] fp32."""
    D = q.size(1) // num_heads
    if scale is None:
        scale = 1.0 / math.sqrt(D)
    return AttnVarlen.apply(q, k, v, cu_seqlens.to(torch.int32), num_heads, scale)
