"""The GPS block's node-level part on the kernel library (hydragnn/ops/node.py: GpsTailPre / GpsTailPost,
reference hydragnn/globalAtt/gps.py:116-150) against the same block written with plain torch ops and
autograd, dropout ON: the Bernoulli masks the native path draws are recorded and replayed in the torch
restatement, so outputs, input / parameter gradients and BatchNorm running statistics are comparable.

  * CPU (`-m "not gpu"`): the autograd Functions over oracle/kernel_spec.py (tests/cpu_patch.py) -- checks
    the backward formulas and the mask bookkeeping;
  * GPU (`-m gpu`): the same comparison with the real kernels, plus every raw entry point
    (hgk_linear_tc, hgk_wgrad_rows, hgk_bn_stats / _apply / _bwd, hgk_mask_axpy) against kernel_spec.
"""
import copy

import pytest
import torch
import torch.nn.functional as F

from oracle import kernel_spec as ks


def _torch_layer(layer, x, h_local, o, masks, scale):
    """gps.py:116-150 with explicit dropout masks (in the order the native path draws them)."""
    it = iter(masks)

    def drop(t):
        m = next(it)
        return t if m is None else t * m.to(t.dtype) * scale

    branches = []
    if h_local is not None:
        branches.append(layer.norm1(drop(h_local) + x))
    h = F.linear(o, layer.attn.out_proj.weight, layer.attn.out_proj.bias)
    branches.append(layer.norm2(drop(h) + x))
    out = sum(branches)
    m = drop(torch.relu(layer.mlp[0](out)))
    m = drop(layer.mlp[3](m))
    return layer.norm3(out + m)


def _run_case(device, with_local, p, N=200, C=128, heads=4, seed=0):
    from hydragnn.globalAtt.gps import GPSConv
    from hydragnn.ops import node

    torch.manual_seed(seed)
    layer = GPSConv(C, None, heads=heads, dropout=p).to(device)
    with torch.no_grad():                      # non-trivial affine parameters and running statistics
        for norm in (layer.norm1, layer.norm2, layer.norm3):
            norm.module.weight.uniform_(0.5, 1.5)
            norm.module.bias.normal_(0, 0.2)
            norm.module.running_mean.normal_(0, 0.3)
            norm.module.running_var.uniform_(0.5, 2.0)
    twin = copy.deepcopy(layer)
    layer.train(), twin.train()
    gen = torch.Generator().manual_seed(seed + 1)
    mk = lambda *s: torch.randn(*s, generator=gen).to(device)
    x, h_local, gy = mk(N, C), (mk(N, C) if with_local else None), mk(N, C)
    leaves = [t.clone().requires_grad_(True) for t in (x, h_local) if t is not None]

    # native: record the masks it draws
    drawn, real = [], node.dropout_mask

    def recording(shape, prob, dev):
        m, s = real(shape, prob, dev)
        drawn.append(m)
        return m, s

    real_many = node.dropout_masks

    def recording_many(shapes, prob, dev):
        ms, s = real_many(shapes, prob, dev)
        drawn.extend(ms)
        return ms, s

    node.dropout_mask = recording
    node.dropout_masks = recording_many
    try:
        xa = leaves[0]
        ha = leaves[1] if with_local else None
        assert layer._native(xa), "native path not selected"
        h1, qkv = layer._tail_pre(xa, ha)
        # stand-in for the attention kernel: any differentiable map qkv -> o
        o = qkv[:, :C] * torch.tanh(qkv[:, C:2 * C]) + qkv[:, 2 * C:]
        y = layer._tail_post(xa, h1, o)
        y.backward(gy)
    finally:
        node.dropout_mask = real
        node.dropout_masks = real_many
    # order drawn: [h_local (pre)], out_proj, mlp hidden, mlp out
    scale = 1.0 / (1.0 - p) if p > 0 else 1.0
    xb = x.clone().requires_grad_(True)
    hb = h_local.clone().requires_grad_(True) if with_local else None
    qkv_b = F.linear(xb, twin.attn.in_proj_weight, twin.attn.in_proj_bias)
    ob = qkv_b[:, :C] * torch.tanh(qkv_b[:, C:2 * C]) + qkv_b[:, 2 * C:]
    yb = _torch_layer(twin, xb, hb, ob, drawn, scale)
    yb.backward(gy)
    return dict(layer=layer, twin=twin, y=y, yb=yb, xa=xa, xb=xb, ha=ha, hb=hb, drawn=drawn)


def _compare(r, rtol, atol):
    close = lambda a, b, what: torch.testing.assert_close(a.detach().cpu(), b.detach().cpu(), rtol=rtol, atol=atol,
                                                          msg=lambda m: f"{what}: {m}")
    close(r["y"], r["yb"], "output")
    close(r["xa"].grad, r["xb"].grad, "d x")
    if r["ha"] is not None:
        close(r["ha"].grad, r["hb"].grad, "d h_local")
    ref = dict(r["twin"].named_parameters())
    for n, prm in r["layer"].named_parameters():
        if ref[n].grad is None:
            assert prm.grad is None, n
            continue
        scale = float(ref[n].grad.abs().max())
        torch.testing.assert_close(prm.grad.cpu(), ref[n].grad.cpu(), rtol=rtol, atol=atol * max(scale, 1.0),
                                   msg=lambda m, n=n: f"grad {n}: {m}")
    refb = dict(r["twin"].named_buffers())
    for n, b in r["layer"].named_buffers():
        torch.testing.assert_close(b.cpu(), refb[n].cpu(), rtol=rtol, atol=atol, msg=lambda m, n=n: f"buffer {n}: {m}")


@pytest.mark.parametrize("with_local,p", [(True, 0.25), (False, 0.25), (True, 0.0)])
def test_native_tail_matches_torch_ops_on_cpu(with_local, p):
    from cpu_patch import cpu_kernels

    with cpu_kernels():
        r = _run_case("cpu", with_local, p)
    assert len(r["drawn"]) == (4 if with_local else 3)
    assert all((m is None) == (p == 0.0) for m in r["drawn"])
    _compare(r, rtol=1e-4, atol=5e-5)     # mlp.3.bias has an analytically zero gradient without dropout: noise vs noise


# ------------------------------------------------------------------------------- GPU ---
@pytest.mark.gpu
@pytest.mark.parametrize("with_local,p,N,C", [(True, 0.25, 200, 128), (False, 0.25, 333, 128), (True, 0.0, 1025, 128),
                                              (True, 0.25, 8192, 128), (True, 0.1, 130, 256)])
def test_native_tail_matches_torch_ops(libhgk, with_local, p, N, C):
    torch.backends.cuda.matmul.allow_tf32 = False
    r = _run_case("cuda:0", with_local, p, N=N, C=C)
    _compare(r, rtol=1e-4, atol=5e-5)


@pytest.mark.gpu
@pytest.mark.parametrize("M,N,K,trans,opts", [
    (1, 128, 128, False, ()), (127, 128, 32, False, ("bias",)), (300, 384, 128, False, ("bias",)),
    (1000, 256, 128, False, ("bias", "relu", "mask")), (513, 128, 256, False, ("bias", "mask", "residual")),
    (700, 256, 128, True, ("nz",)), (8192, 128, 384, True, ("residual",)), (129, 128, 128, True, ()),
    (260, 96, 64, False, ("bias", "residual")), (8192, 384, 128, False, ("bias",)),
    # long K: many wraps of the four-slot TMA ring, accumulation chains of 256 k
    (200, 128, 1024, False, ("bias",)), (333, 256, 1024, True, ())])
def test_linear_tc(libhgk, M, N, K, trans, opts):
    """hgk_linear_tc vs the torch ops (fp32 and, as the truth, fp64)."""
    from hydragnn.ops import node
    import parity_util as pu

    gen = torch.Generator().manual_seed(M + N + K)
    x = torch.randn(M, K, generator=gen)
    w = torch.randn((K, N) if trans else (N, K), generator=gen) / K ** 0.5
    kw = {}
    if "bias" in opts:
        kw["bias"] = torch.randn(N, generator=gen)
    if "relu" in opts:
        kw["relu"] = True
    if "mask" in opts:
        kw["mask"] = (torch.rand(M, N, generator=gen) < 0.75).to(torch.uint8)
        kw["scale"] = 1.0 / 0.75
    if "nz" in opts:
        kw["nz"] = torch.relu(torch.randn(M, N, generator=gen))
        kw["scale"] = 1.25
    if "residual" in opts:
        kw["residual"] = torch.randn(M, N, generator=gen)
    dev = lambda t: t.cuda() if torch.is_tensor(t) else t
    got = node.linear_tc(dev(x), dev(w), trans_w=trans, **{k: dev(v) for k, v in kw.items()})
    ref32 = ks.linear_tc(x, w, trans_w=trans, **kw)
    dbl = lambda t: t.double() if torch.is_tensor(t) and t.is_floating_point() else t
    truth = ks.linear_tc(x.double(), w.double(), trans_w=trans, **{k: dbl(v) for k, v in kw.items()})
    torch.testing.assert_close(got.cpu(), ref32, rtol=1e-5, atol=1e-5)
    pu.assert_close_to_truth(got, ref32, truth, "linear_tc")


@pytest.mark.gpu
def test_linear_tc_strided_operands(libhgk):
    """x / residual as column slices of wider matrices (packed qkv, PaiNN splits): row strides, no copies."""
    from hydragnn.ops import node

    gen = torch.Generator().manual_seed(3)
    wide = torch.randn(500, 384, generator=gen).cuda()
    w = (torch.randn(128, 128, generator=gen) / 11.0).cuda()
    got = node.linear_tc(wide[:, 128:256], w, residual=wide[:, 256:])
    ref = wide[:, 128:256].cpu() @ w.cpu().t() + wide[:, 256:].cpu()
    torch.testing.assert_close(got.cpu(), ref, rtol=1e-5, atol=1e-5)


@pytest.mark.gpu
@pytest.mark.parametrize("M,NA,NB", [(1, 128, 128), (100, 128, 128), (8192, 384, 128), (5000, 128, 256), (777, 256, 256)])
def test_wgrad_rows(libhgk, M, NA, NB):
    from hydragnn.ops import node
    import parity_util as pu

    gen = torch.Generator().manual_seed(M + NA)
    a, b = torch.randn(M, NA, generator=gen), torch.randn(M, NB, generator=gen)
    dw, db = node.wgrad_rows(a.cuda(), b.cuda())
    rw, rb = ks.wgrad_rows(a, b)
    tw, tb = ks.wgrad_rows(a.double(), b.double())
    pu.assert_close_to_truth(dw, rw, tw, "dW")
    pu.assert_close_to_truth(db, rb, tb, "db")
    dw2, db2 = node.wgrad_rows(a.cuda(), b.cuda())
    assert torch.equal(dw, dw2) and torch.equal(db, db2)          # fixed summation order


@pytest.mark.gpu
@pytest.mark.parametrize("N,C", [(2, 128), (127, 128), (129, 64), (8192, 128), (3000, 256)])
def test_batchnorm_kernels(libhgk, N, C):
    """hgk_bn_stats / hgk_bn_apply / hgk_bn_bwd vs nn.BatchNorm1d semantics (kernel_spec) and torch autograd."""
    from hydragnn.ops import node

    gen = torch.Generator().manual_seed(N + C)
    ta = torch.randn(N, C, generator=gen) * 2.0 + 0.7
    tb = torch.randn(N, C, generator=gen) * 0.5 - 1.0
    g = torch.randn(N, C, generator=gen)
    gam = [torch.rand(C, generator=gen) + 0.5 for _ in range(2)]
    bet = [torch.randn(C, generator=gen) for _ in range(2)]
    mask = (torch.rand(N, C, generator=gen) < 0.8).to(torch.uint8)
    rm, rv = torch.randn(C, generator=gen), torch.rand(C, generator=gen) + 0.5
    nbt = torch.tensor(3, dtype=torch.int64)
    rm_d, rv_d, nbt_d = rm.clone().cuda(), rv.clone().cuda(), nbt.clone().cuda()

    mean, rstd = node.bn_stats(ta.cuda(), rm_d, rv_d, nbt_d, 0.1, 1e-5)
    rm_r, rv_r, nbt_r = rm.clone(), rv.clone(), nbt.clone()
    mean_r, rstd_r = ks.bn_stats(ta, rm_r, rv_r, nbt_r, 0.1, 1e-5)
    tol = dict(rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(mean.cpu(), mean_r, **tol)
    torch.testing.assert_close(rstd.cpu(), rstd_r, **tol)
    torch.testing.assert_close(rm_d.cpu(), rm_r, **tol)
    torch.testing.assert_close(rv_d.cpu(), rv_r, **tol)
    assert int(nbt_d) == 4
    mean_b, rstd_b = node.bn_stats(tb.cuda())

    pa = node.BnParams(mean, rstd, gam[0].cuda(), bet[0].cuda())
    pb = node.BnParams(mean_b, rstd_b, gam[1].cuda(), bet[1].cuda())
    # torch autograd of the same function of (ta, tb, gamma, beta): batch statistics inside the graph
    leaves = [t.clone().requires_grad_(True) for t in (ta, tb, gam[0], bet[0], gam[1], bet[1])]
    ya = F.batch_norm(leaves[0], None, None, leaves[2], leaves[3], True, 0.1, 1e-5)
    yb = F.batch_norm(leaves[1], None, None, leaves[4], leaves[5], True, 0.1, 1e-5)
    torch.testing.assert_close(node.bn_apply(ta.cuda(), pa).cpu(), ya.detach(), rtol=1e-5, atol=2e-6)
    both = node.bn_apply(ta.cuda(), pa, tb.cuda(), pb, relu=True)
    torch.testing.assert_close(both.cpu(), torch.relu(ya + yb).detach(), rtol=1e-5, atol=2e-6)
    (ya + yb).backward(g)
    r = node.bn_bwd(g.cuda(), ta.cuda(), pa, tb.cuda(), pb, mask_b=mask.cuda(), scale_b=1.25)
    gtol = dict(rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(r["dta"].cpu(), leaves[0].grad, **gtol)
    torch.testing.assert_close(r["dtb"].cpu(), leaves[1].grad, **gtol)
    torch.testing.assert_close(r["dzb"].cpu(), leaves[1].grad * mask * 1.25, **gtol)
    assert r["dza"] is r["dta"]
    for key, leaf in (("dgamma_a", leaves[2]), ("dbeta_a", leaves[3]), ("dgamma_b", leaves[4]), ("dbeta_b", leaves[5])):
        torch.testing.assert_close(r[key].cpu(), leaf.grad, rtol=1e-4, atol=1e-4 * max(1.0, float(leaf.grad.abs().max())))
    single = node.bn_bwd(g.cuda(), ta.cuda(), pa, mask_a=mask.cuda(), scale_a=2.0)
    torch.testing.assert_close(single["dza"].cpu(), leaves[0].grad * mask * 2.0, **gtol)
    z = node.mask_axpy(ta.cuda(), mask.cuda(), 1.25, tb.cuda())
    torch.testing.assert_close(z.cpu(), ta * mask * 1.25 + tb, rtol=1e-6, atol=1e-6)


@pytest.mark.gpu
def test_linear_fast_autograd(libhgk):
    """LinearTC (first and, under create_graph, second derivatives) vs F.linear."""
    from hydragnn.ops import node

    gen = torch.Generator().manual_seed(8)
    x, w, b = torch.randn(300, 128, generator=gen), torch.randn(256, 128, generator=gen) / 11, torch.randn(256, generator=gen)
    g = torch.randn(300, 256, generator=gen)
    cpu = [t.clone().requires_grad_(True) for t in (x, w, b)]
    F.linear(*cpu).backward(g)
    gpu = [t.clone().cuda().requires_grad_(True) for t in (x, w, b)]
    node.linear_fast(*gpu).backward(g.cuda())
    for a, c in zip(gpu, cpu):
        torch.testing.assert_close(a.grad.cpu(), c.grad, rtol=1e-4, atol=1e-4)
    # double backward (the force loss differentiates the backward pass)
    xs = x.clone().cuda().requires_grad_(True)
    y = node.linear_fast(xs, gpu[1].detach().requires_grad_(True), None)
    (gx,) = torch.autograd.grad(y.pow(2).sum(), xs, create_graph=True)
    gx.sum().backward()
    xc = x.clone().requires_grad_(True)
    yc = F.linear(xc, w)
    (gxc,) = torch.autograd.grad(yc.pow(2).sum(), xc, create_graph=True)
    gxc.sum().backward()
    torch.testing.assert_close(xs.grad.cpu(), xc.grad, rtol=1e-4, atol=1e-3)


@pytest.mark.gpu
@pytest.mark.parametrize("N,nv,C", [(192, 187, 128), (512, 300, 128), (256, 256, 64), (1030, 5, 128)])
def test_batchnorm_valid_rows(libhgk, N, nv, C):
    """n_valid (device scalar): statistics and gradients over the first nv rows only; rows beyond get no
    gradient -- the padded mini-batches of train/step_graph.py."""
    from hydragnn.ops import node

    gen = torch.Generator().manual_seed(N + nv)
    t = torch.randn(N, C, generator=gen) * 2 + 0.5
    t[nv:] = 1.0e3                                                  # pad rows must not leak into the statistics
    g = torch.randn(N, C, generator=gen)
    gam, bet = torch.rand(C, generator=gen) + 0.5, torch.randn(C, generator=gen)
    nv_d = torch.tensor([nv], dtype=torch.int64, device="cuda")
    mean, rstd = node.bn_stats(t.cuda(), n_valid=nv_d)
    mean_r, rstd_r = ks.bn_stats(t, n_valid=nv)
    torch.testing.assert_close(mean.cpu(), mean_r, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(rstd.cpu(), rstd_r, rtol=1e-5, atol=1e-6)
    pa = node.BnParams(mean, rstd, gam.cuda(), bet.cuda())
    pr = node.BnParams(mean_r, rstd_r, gam, bet)
    got = node.bn_bwd(g.cuda(), t.cuda(), pa, n_valid=nv_d)
    want = ks.bn_bwd(g, t, pr, n_valid=nv)
    torch.testing.assert_close(got["dta"].cpu()[:nv], want["dta"][:nv], rtol=1e-4, atol=1e-5)
    assert float(got["dta"][nv:].abs().max()) == 0.0 if nv < N else True
    torch.testing.assert_close(got["dgamma_a"].cpu(), want["dgamma_a"], rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(got["dbeta_a"].cpu(), want["dbeta_a"], rtol=1e-4, atol=1e-4)


@pytest.mark.gpu
def test_attention_pad_segment_is_invisible(libhgk):
    """cu_seqlens = [0, N, N_b]: the rows of the first segment equal whole-batch attention over N rows."""
    from hydragnn.ops import attn_varlen

    gen = torch.Generator().manual_seed(1)
    N, Nb, H, D = 187, 192, 4, 32
    qkv = torch.randn(Nb, 3 * H * D, generator=gen).cuda()
    c = H * D
    whole = attn_varlen(qkv[:N, :c], qkv[:N, c:2 * c], qkv[:N, 2 * c:], torch.tensor([0, N], dtype=torch.int32).cuda(), H)
    padded = attn_varlen(qkv[:, :c], qkv[:, c:2 * c], qkv[:, 2 * c:], torch.tensor([0, N, Nb], dtype=torch.int32).cuda(), H)
    torch.testing.assert_close(padded[:N], whole, rtol=1e-6, atol=1e-6)
