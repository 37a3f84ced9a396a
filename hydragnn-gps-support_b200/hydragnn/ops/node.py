"""Node-row operators of the GPS block on libhgk: tensor-core Linears with fused epilogues,
BatchNorm with batch statistics, and the two autograd Functions that make up the layer's
node-level part around the attention kernel.

Reference: hydragnn/globalAtt/gps.py:72-83,116-150 --

    h1  = norm1(dropout(conv(x)) + x)
    h2  = norm2(dropout(out_proj(attention(in_proj(x)))) + x)
    out = h1 + h2;  out = norm3(out + dropout(W2 dropout(relu(W1 out + b1)) + b2))

run there as ~65 torch launches per layer and step (cuBLAS fp32 GEMMs, four BatchNorm kernels per
norm, dropout, adds).  Here: `hgk_linear_tc` (tcgen05, 3xTF32: fp32-accurate) with bias / ReLU /
dropout mask / residual applied in the accumulator read-out, `hgk_bn_stats` + `hgk_bn_apply`
(two-pass statistics per slab, fixed merge order), `hgk_bn_bwd`, and `hgk_wgrad_rows` for the weight
and bias gradients.  Dropout masks are torch's own Bernoulli draws (uint8), so the random stream is
torch's; forward and backward read the same mask.
"""
import os

import torch
import torch.nn.functional as F

from .. import _hgk


# ------------------------------------------------------------------------- raw calls ---
def supported_width(*widths):
    return all(w % 128 == 0 for w in widths)


def device_ok(t):
    """The kernels take fp32 CUDA tensors (tests/cpu_patch.py swaps this together with the launchers)."""
    return t.is_cuda and t.dtype == torch.float32


def _rows(t):
    """[M, K] fp32 with unit column stride and 16-byte aligned rows -> (tensor, ld)."""
    if t.dtype != torch.float32 or t.dim() != 2:
        raise _hgk.HgkError(f"expected a 2-D float32 tensor, got {t.dtype} {tuple(t.shape)}")
    if t.stride(1) != 1 or t.stride(0) % 4 or t.data_ptr() % 16 or (t.size(0) > 1 and t.stride(0) < t.size(1)):
        t = t.contiguous()
    return t, (t.stride(0) if t.size(0) > 1 else max(t.size(1), 4))


def linear_tc(x, w, bias=None, trans_w=False, relu=False, mask=None, nz=None, scale=1.0, residual=None):
    """y = residual + scale * mask * [nz != 0] * act(x w^T + bias)   (trans_w: w is [K, N], y = x w)."""
    x, ldx = _rows(x)
    w, ldw = _rows(w)
    M, K = x.shape
    N = w.size(1) if trans_w else w.size(0)
    y = torch.empty((M, N), dtype=torch.float32, device=x.device)
    nz_t, ldnz = (None, 0) if nz is None else _rows(nz)
    r_t, ldr = (None, 0) if residual is None else _rows(residual)
    if mask is not None and not mask.is_contiguous():
        mask = mask.contiguous()
    if bias is not None and not bias.is_contiguous():
        bias = bias.contiguous()
    nbytes = 4 * (M * K + N * K + M * N * (1 + (residual is not None) + (nz is not None))) + (M * N if mask is not None else 0)
    _hgk.call("hgk_linear_tc", _hgk.ptr(y), N, _hgk.ptr(x), ldx, _hgk.ptr(w), ldw, int(bool(trans_w)), _hgk.ptr(bias),
              int(bool(relu)), _hgk.ptr(mask), _hgk.ptr(nz_t), ldnz, float(scale), _hgk.ptr(r_t), ldr, M, N, K,
              cost=(nbytes, 2 * M * N * K))
    return y


def wgrad_rows(a, b, want_bias=True):
    """(dW [NA, NB], db [NA] or None) = (a^T b, column sums of a) for row matrices a [M, NA], b [M, NB]."""
    a, lda = _rows(a)
    b, ldb = _rows(b)
    M, NA = a.shape
    NB = b.size(1)
    L = _hgk.lib()
    ws = torch.empty(L.hgk_wgrad_rows_workspace_floats(M, NA, NB), dtype=torch.float32, device=a.device)
    dw = torch.empty((NA, NB), dtype=torch.float32, device=a.device)
    db = torch.empty(NA, dtype=torch.float32, device=a.device) if want_bias else None
    _hgk.call("hgk_wgrad_rows", _hgk.ptr(dw), NB, _hgk.ptr(db), _hgk.ptr(ws), _hgk.ptr(a), lda, NA, _hgk.ptr(b), ldb, NB, M,
              cost=(4 * (M * (NA + NB) + 2 * ws.numel()), 2 * M * NA * NB))
    return dw, db


class BnParams:
    """What the kernels need of one BatchNorm: statistics of this batch + the affine parameters."""

    __slots__ = ("mean", "rstd", "gamma", "beta")

    def __init__(self, mean, rstd, gamma, beta):
        self.mean, self.rstd, self.gamma, self.beta = mean, rstd, gamma, beta


# Rows of the node matrices that belong to the mini-batch proper: None (all) or a device int64 scalar set by
# train/step_graph.py around the forward pass of a batch padded to its size bucket.
valid_rows = None


def bn_stats(t, running_mean=None, running_var=None, num_batches_tracked=None, momentum=0.1, eps=1e-5, n_valid=None):
    """Batch statistics of the columns of t [N, C] -> (mean, 1 / sqrt(var + eps)); updates the running
    buffers in place the way nn.BatchNorm1d does in training (unbiased variance, momentum)."""
    t, ldt = _rows(t)
    N, C = t.shape
    L = _hgk.lib()
    S = L.hgk_bn_slabs(N)
    mean = torch.empty(C, dtype=torch.float32, device=t.device)
    rstd = torch.empty(C, dtype=torch.float32, device=t.device)
    partial = torch.empty(2 * S * C, dtype=torch.float32, device=t.device)
    _hgk.call("hgk_bn_stats", _hgk.ptr(mean), _hgk.ptr(rstd), _hgk.ptr(running_mean), _hgk.ptr(running_var),
              _hgk.ptr(num_batches_tracked), float(momentum), float(eps), _hgk.ptr(partial), _hgk.ptr(t), ldt, N, C,
              _hgk.ptr(n_valid), cost=(4 * (2 * N * C + 4 * S * C), 0))
    return mean, rstd


def bn_apply(ta, pa, tb=None, pb=None, relu=False):
    """act( BN_a(ta) [+ BN_b(tb)] ) with the given statistics / affine parameters."""
    ta, lda = _rows(ta)
    N, C = ta.shape
    out = torch.empty((N, C), dtype=torch.float32, device=ta.device)
    if tb is not None:
        tb, ldb = _rows(tb)
        side_b = (_hgk.ptr(tb), ldb, _hgk.ptr(pb.mean), _hgk.ptr(pb.rstd), _hgk.ptr(pb.gamma), _hgk.ptr(pb.beta))
    else:
        side_b = (_hgk.ptr(None), 0, _hgk.ptr(None), _hgk.ptr(None), _hgk.ptr(None), _hgk.ptr(None))
    _hgk.call("hgk_bn_apply", _hgk.ptr(out), C, _hgk.ptr(ta), lda, _hgk.ptr(pa.mean), _hgk.ptr(pa.rstd),
              _hgk.ptr(pa.gamma), _hgk.ptr(pa.beta), *side_b, int(bool(relu)), N, C,
              cost=(4 * N * C * (2 + (tb is not None)), 0))
    return out


def bn_bwd(g, ta, pa, tb=None, pb=None, mask_a=None, scale_a=1.0, mask_b=None, scale_b=1.0, n_valid=None):
    """Backward of bn_apply (without the activation) for the gradient g of its output.
    -> dict(dta, dza, dgamma_a, dbeta_a[, dtb, dzb, dgamma_b, dbeta_b]); dz* = scale * mask * dt*
    (the gradient behind a dropout feeding the BatchNorm input) or dt* itself without a mask."""
    g, ldg = _rows(g)
    ta, lda = _rows(ta)
    N, C = ta.shape
    dev = ta.device
    new = lambda *shape: torch.empty(shape, dtype=torch.float32, device=dev)
    L = _hgk.lib()
    ws = new(L.hgk_bn_bwd_workspace_floats(N, C))
    out = dict(dta=new(N, C), dgamma_a=new(C) if pa.gamma is not None else None, dbeta_a=new(C))
    out["dza"] = new(N, C) if mask_a is not None else None
    two = tb is not None
    if two:
        tb, ldb = _rows(tb)
        out.update(dtb=new(N, C), dgamma_b=new(C) if pb.gamma is not None else None, dbeta_b=new(C))
        out["dzb"] = new(N, C) if mask_b is not None else None
        side_b = (_hgk.ptr(tb), ldb, _hgk.ptr(pb.mean), _hgk.ptr(pb.rstd), _hgk.ptr(pb.gamma))
    else:
        ldb = 0
        side_b = (_hgk.ptr(None), 0, _hgk.ptr(None), _hgk.ptr(None), _hgk.ptr(None))
    _hgk.call("hgk_bn_bwd", _hgk.ptr(out["dta"]), _hgk.ptr(out["dza"]), _hgk.ptr(mask_a), float(scale_a),
              _hgk.ptr(out["dgamma_a"]), _hgk.ptr(out["dbeta_a"]),
              _hgk.ptr(out.get("dtb")), _hgk.ptr(out.get("dzb")), _hgk.ptr(mask_b), float(scale_b),
              _hgk.ptr(out.get("dgamma_b")), _hgk.ptr(out.get("dbeta_b")), _hgk.ptr(ws), _hgk.ptr(g), ldg,
              _hgk.ptr(ta), lda, _hgk.ptr(pa.mean), _hgk.ptr(pa.rstd), _hgk.ptr(pa.gamma), *side_b, N, C,
              _hgk.ptr(n_valid), cost=(4 * N * C * (3 + 2 * (1 + two) + (mask_a is not None) + (mask_b is not None)), 0))
    if out["dza"] is None:
        out["dza"] = out["dta"]
    if two and out["dzb"] is None:
        out["dzb"] = out["dtb"]
    return out


def mask_axpy(z, mask=None, scale=1.0, residual=None):
    """residual + scale * mask * z."""
    z, ldz = _rows(z)
    N, C = z.shape
    out = torch.empty((N, C), dtype=torch.float32, device=z.device)
    r_t, ldr = (None, 0) if residual is None else _rows(residual)
    _hgk.call("hgk_mask_axpy", _hgk.ptr(out), _hgk.ptr(z), ldz, _hgk.ptr(mask), float(scale), _hgk.ptr(r_t), ldr, N, C,
              cost=(4 * N * C * (2 + (residual is not None)) + (N * C if mask is not None else 0), 0))
    return out


def dropout_mask(shape, p, device):
    """(uint8 keep mask, 1 / (1 - p)) drawn from torch's CUDA generator; (None, 1.0) when p == 0."""
    if p <= 0.0:
        return None, 1.0
    keep = torch.empty(shape, dtype=torch.uint8, device=device).bernoulli_(1.0 - p)
    return keep, 1.0 / (1.0 - p)


def dropout_masks(shapes, p, device):
    """Several keep masks from ONE Bernoulli launch (each a contiguous chunk of one buffer) + the common scale."""
    if p <= 0.0:
        return [None] * len(shapes), 1.0
    sizes = [int(a) * int(b) for a, b in shapes]
    # chunk starts kept 16-byte aligned (the kernels read the masks with 16-byte loads)
    starts, total = [], 0
    for n in sizes:
        starts.append(total)
        total += (n + 15) // 16 * 16
    buf = torch.empty(total, dtype=torch.uint8, device=device).bernoulli_(1.0 - p)
    return [buf[st:st + n].view(shape) for st, n, shape in zip(starts, sizes, shapes)], 1.0 / (1.0 - p)


# ------------------------------------------------------------------ autograd: Linear ---
# Weight gradients on a side stream.  In the backward pass of a Linear the weight / bias gradient (hgk_wgrad_rows)
# and the input gradient (hgk_linear_tc) are independent and both are small grids (64-192 CTAs on 148 SMs): run side
# by side they fill the GPU.  Done only while a CUDA graph is being captured (train/step_graph.py, graphed_tail.py),
# where the fork and the join are graph edges and cost nothing at replay; eagerly the event calls would cost more
# host time than the overlap saves.  HYDRAGNN_SIDE_WGRAD=0 switches it off.
_SIDE_WGRAD = os.environ.get("HYDRAGNN_SIDE_WGRAD", "1") != "0"
_side_streams = {}


class _SideWgrad:
    """`with side: ...` runs the enclosed launches on the side stream (which first waits for the work already on
    the current stream); `side.join()` makes the current stream wait for them.  Inactive outside a capture."""

    def __init__(self, tensor):
        self.active = bool(_SIDE_WGRAD and tensor.is_cuda and torch.cuda.is_current_stream_capturing())
        self.used = False
        if self.active:
            dev = tensor.device
            self.cur = torch.cuda.current_stream(dev)
            if dev not in _side_streams:
                _side_streams[dev] = torch.cuda.Stream(dev)
            self.side = _side_streams[dev]
            self.ctx = None

    def __enter__(self):
        if self.active:
            self.side.wait_stream(self.cur)
            self.ctx = torch.cuda.stream(self.side)
            self.ctx.__enter__()
            self.used = True
        return self

    def __exit__(self, *exc):
        if self.active:
            self.ctx.__exit__(*exc)
        return False

    def join(self):
        if self.active and self.used:
            self.cur.wait_stream(self.side)


class LinearTC(torch.autograd.Function):
    """y = x W^T + b on the tensor cores.  Backward: input gradient by the same kernel on the
    transposed weight, weight / bias gradients by hgk_wgrad_rows; when the backward pass is itself
    being differentiated (create_graph=True, force loss) it is written with differentiable torch ops."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        ctx.save_for_backward(x, weight)
        ctx.has_bias = bias is not None
        return linear_tc(x, weight, bias)

    @staticmethod
    def backward(ctx, g):
        x, weight = ctx.saved_tensors
        if torch.is_grad_enabled():
            gx = g @ weight if ctx.needs_input_grad[0] else None
            gw = g.t() @ x if ctx.needs_input_grad[1] else None
            gb = g.sum(0) if ctx.has_bias and ctx.needs_input_grad[2] else None
            return gx, gw, gb
        side = _SideWgrad(g)
        gw = gb = None
        if ctx.needs_input_grad[1] or (ctx.has_bias and ctx.needs_input_grad[2]):
            with side:
                gw, gb = wgrad_rows(g, x, want_bias=ctx.has_bias)
        gx = linear_tc(g, weight, trans_w=True) if ctx.needs_input_grad[0] else None
        side.join()
        return gx, gw, gb


def linear_fast(x, weight, bias=None):
    """F.linear for node rows: the tensor-core kernel when the shapes allow it (both widths multiples
    of 128, fp32 CUDA rows), torch otherwise."""
    if (device_ok(x) and x.dim() == 2 and weight.dtype == torch.float32
            and supported_width(weight.size(0), weight.size(1)) and x.size(0) > 0):
        return LinearTC.apply(x, weight, bias)
    return F.linear(x, weight, bias)


# --------------------------------------------------------- autograd: the GPS block tail ---
def _bn_buffers(norm):
    bn = norm.module
    return bn.running_mean, bn.running_var, bn.num_batches_tracked, bn.momentum, bn.eps


class GpsTailPre(torch.autograd.Function):
    """(x, h_local) -> (t1 = dropout(h_local) + x  [the input of norm1], qkv = in_proj(x))."""

    @staticmethod
    def forward(ctx, x, h_local, in_w, in_b, p):
        mask, scale = (None, 1.0)
        t1 = None
        if h_local is not None:
            mask, scale = dropout_mask(h_local.shape, p, x.device)
            t1 = mask_axpy(h_local, mask, scale, x)
        qkv = linear_tc(x, in_w, in_b)
        ctx.save_for_backward(x, in_w, mask)
        ctx.scale, ctx.has_local = scale, h_local is not None
        if t1 is None:
            return qkv
        return t1, qkv

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, *grads):
        x, in_w, mask = ctx.saved_tensors
        dt1, dqkv = grads if ctx.has_local else (None, grads[0])
        side = _SideWgrad(dqkv)
        with side:
            dw, db = wgrad_rows(dqkv, x)
        dx = linear_tc(dqkv, in_w, trans_w=True, residual=dt1)
        dh = None
        if ctx.has_local:
            dh = mask_axpy(dt1, mask, ctx.scale) if mask is not None else dt1
        side.join()
        return dx, dh, dw, db, None


class GpsTailPost(torch.autograd.Function):
    """(x, t1, o) -> norm3(s + dropout(W2 dropout(relu(W1 s + b1)) + b2)),
    s = norm1(t1) + norm2(dropout(out_proj(o)) + x)   (batch statistics)."""

    @staticmethod
    def forward(ctx, x, t1, o, out_w, out_b, w1, b1, w2, b2, g1, be1, g2, be2, g3, be3, bufs, p):
        dev = x.device
        N, C = x.shape
        nv = valid_rows                     # device scalar for a padded batch (train/step_graph.py), else None
        (m2, m3, m4), scale = dropout_masks([(N, C), (N, w1.size(0)), (N, C)], p, dev)   # one launch for the three
        t2 = linear_tc(o, out_w, out_b, mask=m2, scale=scale, residual=x)
        p2 = BnParams(*bn_stats(t2, *bufs[1], n_valid=nv), g2, be2)
        if t1 is not None:
            p1 = BnParams(*bn_stats(t1, *bufs[0], n_valid=nv), g1, be1)
            s = bn_apply(t1, p1, t2, p2)
        else:
            p1 = None
            s = bn_apply(t2, p2)
        u = linear_tc(s, w1, b1, relu=True, mask=m3, scale=scale)
        t3 = linear_tc(u, w2, b2, mask=m4, scale=scale, residual=s)
        p3 = BnParams(*bn_stats(t3, *bufs[2], n_valid=nv), g3, be3)
        y = bn_apply(t3, p3)
        ctx.save_for_backward(o, t1, t2, t3, s, u, m2, m4, out_w, w1, w2, g1, g2, g3,
                              p1.mean if p1 else None, p1.rstd if p1 else None, p2.mean, p2.rstd, p3.mean, p3.rstd)
        ctx.scale, ctx.nv = scale, nv
        return y

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, gy):
        (o, t1, t2, t3, s, u, m2, m4, out_w, w1, w2, g1, g2, g3, mean1, rstd1, mean2, rstd2, mean3, rstd3) = ctx.saved_tensors
        scale, nv = ctx.scale, ctx.nv
        r3 = bn_bwd(gy, t3, BnParams(mean3, rstd3, g3, None), mask_a=m4, scale_a=scale, n_valid=nv)
        dt3, dz4 = r3["dta"], r3["dza"]                       # dz4: gradient at W2's output (behind its dropout)
        side = _SideWgrad(gy)                                 # the three weight gradients run beside the main chain
        with side:
            dw2, db2 = wgrad_rows(dz4, u)
        dzu = linear_tc(dz4, w2, trans_w=True, nz=u, scale=scale)      # through dropout and ReLU: where u != 0
        with side:
            dw1, db1 = wgrad_rows(dzu, s)
        ds = linear_tc(dzu, w1, trans_w=True, residual=dt3)
        p2 = BnParams(mean2, rstd2, g2, None)
        if t1 is not None:
            r = bn_bwd(ds, t1, BnParams(mean1, rstd1, g1, None), t2, p2, mask_b=m2, scale_b=scale, n_valid=nv)
            dt1, dt2, dz2 = r["dta"], r["dtb"], r["dzb"]
            dg1, dbe1, dg2, dbe2 = r["dgamma_a"], r["dbeta_a"], r["dgamma_b"], r["dbeta_b"]
        else:
            r = bn_bwd(ds, t2, p2, mask_a=m2, scale_a=scale, n_valid=nv)
            dt1, dt2, dz2 = None, r["dta"], r["dza"]
            dg1 = dbe1 = None
            dg2, dbe2 = r["dgamma_a"], r["dbeta_a"]
        with side:
            dwo, dbo = wgrad_rows(dz2, o)
        do = linear_tc(dz2, out_w, trans_w=True)
        side.join()
        return (dt2, dt1, do, dwo, dbo, dw1, db1, dw2, db2, dg1, dbe1, dg2, dbe2, r3["dgamma_a"], r3["dbeta_a"],
                None, None)
