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


def _attn_bwd_raw(do, o, lse, q, k, v, cu, H, scale, packed=False):
    """dq, dk, dv [N, H*D]; packed=True: the three column blocks of ONE [N, 3*H*D] tensor, which is returned
    instead (the gradient of a packed in-projection output: no slice-backward / accumulate pass afterwards)."""
    N, HD = o.shape
    D = HD // H
    do_t, lddo = _hgk.rows(do)
    if packed:
        dqkv = torch.empty((N, 3 * HD), dtype=torch.float32, device=o.device)
        dq, dk, dv = dqkv[:, :HD], dqkv[:, HD:2 * HD], dqkv[:, 2 * HD:]
        ldg = 3 * HD
    else:
        dq = torch.empty((N, HD), dtype=torch.float32, device=o.device)
        dk = torch.empty_like(dq)
        dv = torch.empty_like(dq)
        ldg = HD
    delta = torch.empty((N, H), dtype=torch.float32, device=o.device)
    q_t, ldq = _hgk.rows(q)
    k_t, ldk = _hgk.rows(k)
    v_t, ldv = _hgk.rows(v)
    _hgk.call("hgk_attn_varlen_bwd_ld", _hgk.ptr(dq), _hgk.ptr(dk), _hgk.ptr(dv), ldg, _hgk.ptr(delta),
              _hgk.ptr(do_t), lddo, _hgk.ptr(o), HD, _hgk.ptr(lse), _hgk.ptr(q_t), ldq, _hgk.ptr(k_t), ldk,
              _hgk.ptr(v_t), ldv, _hgk.ptr(cu), cu.numel() - 1, N, H, D, float(scale),
              cost=(4 * (8 * N * HD + 2 * N * H), 10 * HD * _sum_sq(cu)))
    return dqkv if packed else (dq, dk, dv)


def _segment_bounds(cu):
    """Segment boundaries as host integers.  Callers that know them attach `_hgk_bounds` to the tensor
    (Base._attention_segments does); otherwise this is a device -> host read."""
    b = getattr(cu, "_hgk_bounds", None)
    if b is None:
        b = [int(c) for c in cu.tolist()]
    return b


# rows of one score block of the differentiable backward: memory is H * _DENSE_BLOCK * L floats per
# temporary instead of H * L * L (1 GB per temporary at H = 4, L = 8192)
_DENSE_BLOCK = 1024


def _dense_block(do_b, q_b, ks, vs, scale):
    """Backward terms of one block of query rows, all heads: (dq_b, dk partial, dv partial).
    q_b, do_b: [H, B, D]; ks, vs: [H, L, D].  Needs the softmax statistics of complete rows only,
    so a block of query rows is self-contained."""
    p = torch.softmax((q_b @ ks.transpose(1, 2)) * scale, dim=-1)          # [H, B, L]
    dp = do_b @ vs.transpose(1, 2)
    ds = p * (dp - (dp * p).sum(dim=-1, keepdim=True))
    return (ds @ ks) * scale, (ds.transpose(1, 2) @ q_b) * scale, p.transpose(1, 2) @ do_b


def _dense_backward(do, q, k, v, cu, H, scale):
    """(dq, dk, dv) from differentiable torch ops.  Only used when the backward pass is itself being
    differentiated (create_graph=True: a force loss on a model whose attention inputs depend on the
    positions) -- the kernels provide first derivatives only.  Works segment by segment and, inside a
    segment, in blocks of `_DENSE_BLOCK` query rows under activation checkpointing: the [H, B, L] score
    blocks are recomputed when the second derivative is taken instead of being kept alive, so memory
    stays O(H * B * L) per block rather than O(H * L^2) per layer."""
    from torch.utils.checkpoint import checkpoint

    N, HD = q.shape
    D = HD // H
    bounds = _segment_bounds(cu)
    dq, dk, dv = [], [], []
    for b, e in zip(bounds[:-1], bounds[1:]):
        L = e - b
        if L == 0:
            continue
        qs, ks, vs, gs = (t[b:e].reshape(L, H, D).transpose(0, 1) for t in (q, k, v, do))      # [H, L, D]
        dq_s, dk_s, dv_s = [], None, None
        for r0 in range(0, L, _DENSE_BLOCK):
            r1 = min(L, r0 + _DENSE_BLOCK)
            if L > _DENSE_BLOCK:
                dq_b, dk_b, dv_b = checkpoint(_dense_block, gs[:, r0:r1], qs[:, r0:r1], ks, vs, scale,
                                              use_reentrant=False)
            else:
                dq_b, dk_b, dv_b = _dense_block(gs[:, r0:r1], qs[:, r0:r1], ks, vs, scale)
            dq_s.append(dq_b)
            dk_s = dk_b if dk_s is None else dk_s + dk_b
            dv_s = dv_b if dv_s is None else dv_s + dv_b
        dq.append(torch.cat(dq_s, dim=1).transpose(0, 1).reshape(L, HD))
        dk.append(dk_s.transpose(0, 1).reshape(L, HD))
        dv.append(dv_s.transpose(0, 1).reshape(L, HD))
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
        ctx.bounds = getattr(cu_seqlens, "_hgk_bounds", None)     # python attributes do not survive save_for_backward
        return o

    @staticmethod
    def backward(ctx, do):
        q, k, v, o, lse, cu = ctx.saved_tensors
        if torch.is_grad_enabled():        # create_graph=True: stay differentiable
            if ctx.bounds is not None:
                cu._hgk_bounds = ctx.bounds
            dq, dk, dv = _dense_backward(do, q, k, v, cu, ctx.H, ctx.scale)
        else:
            dq, dk, dv = _attn_bwd_raw(do, o, lse, q, k, v, cu, ctx.H, ctx.scale)
        return dq, dk, dv, None, None, None


class AttnVarlenPacked(torch.autograd.Function):
    """The same on the packed in-projection output qkv [N, 3*H*D] (q | k | v column blocks): ONE autograd
    edge carries the [N, 3*H*D] gradient, written in place by the backward kernels."""

    @staticmethod
    def forward(ctx, qkv, cu_seqlens, H, scale):
        c = qkv.size(1) // 3
        o, lse, *_ = _attn_fwd_raw(qkv[:, :c], qkv[:, c:2 * c], qkv[:, 2 * c:], cu_seqlens, H, scale)
        ctx.save_for_backward(qkv, o, lse, cu_seqlens)
        ctx.H, ctx.scale = H, scale
        ctx.bounds = getattr(cu_seqlens, "_hgk_bounds", None)
        return o

    @staticmethod
    def backward(ctx, do):
        qkv, o, lse, cu = ctx.saved_tensors
        c = qkv.size(1) // 3
        q, k, v = qkv[:, :c], qkv[:, c:2 * c], qkv[:, 2 * c:]
        if torch.is_grad_enabled():        # create_graph=True: stay differentiable
            if ctx.bounds is not None:
                cu._hgk_bounds = ctx.bounds
            return torch.cat(_dense_backward(do, q, k, v, cu, ctx.H, ctx.scale), dim=1), None, None, None
        return _attn_bwd_raw(do, o, lse, q, k, v, cu, ctx.H, ctx.scale, packed=True), None, None, None


def attn_varlen_packed(qkv, cu_seqlens, num_heads, scale=None):
    """attn_varlen(qkv[:, :C], qkv[:, C:2C], qkv[:, 2C:]) for a packed [N, 3C] projection output."""
    D = qkv.size(1) // (3 * num_heads)
    if scale is None:
        scale = 1.0 / math.sqrt(D)
    return AttnVarlenPacked.apply(qkv, cu_seqlens.to(torch.int32), num_heads, scale)


def attn_varlen(q, k, v, cu_seqlens, num_heads, scale=None):
    """softmax(q k^T * scale) v inside each segment; q,k,v [N, H*D] fp32."""
    D = q.size(1) // num_heads
    if scale is None:
        scale = 1.0 / math.sqrt(D)
    return AttnVarlen.apply(q, k, v, cu_seqlens.to(torch.int32), num_heads, scale)
