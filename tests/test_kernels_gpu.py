"""`-m gpu`: every entry point of include/hgk.h, called through the C ABI (ctypes), against its
CPU restatement in oracle/kernel_spec.py on the same seeded inputs.

Bars: bit-exact for index work (CSR, edge lists) and for the "segment order" reductions hgk.h
promises (contract_node, contract_edge_w, edge_gather_add); stated fp32 tolerances for the tree
reductions (contract_edge_u), PNA statistics and attention."""
import numpy as np
import pytest
import torch

from oracle import kernel_spec as ks
import parity_util as pu

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def ops(libhgk):
    from hydragnn import ops as _ops
    from hydragnn.ops import attention, contract, pna, topo

    return dict(ops=_ops, contract=contract, pna=pna, attention=attention, topo=topo)


def random_graph(gen, N, E, sorted_dst=False):
    src = torch.randint(0, N, (E,), generator=gen)
    dst = torch.randint(0, N, (E,), generator=gen)
    if sorted_dst:
        dst = torch.sort(dst).values
    return torch.stack([src, dst])


def make_topo(ops, ei, N, sorted_side=None):
    return ops["topo"].Topo.from_edge_index(ei.to(DEV), N, sorted_side=sorted_side)


# ----------------------------------------------------------------------------- CSR ---
@pytest.mark.parametrize("N,E", [(1, 1), (7, 0), (50, 400), (1000, 20000), (8192, 475000), (3, 5000),
                                 # long rows: one sorted chunk (<= 4096) and several chunks merged by rank
                                 (2, 300), (3, 20000), (1, 9000)])
def test_csr_build_bit_exact(ops, N, E):
    gen = torch.Generator().manual_seed(N + E)
    key = torch.randint(0, N, (E,), generator=gen, dtype=torch.int32)
    rowptr, perm = ops["topo"].build_csr(key.to(DEV), N)
    r_ref, p_ref = ks.csr_build(key, N)
    assert torch.equal(rowptr.cpu(), r_ref)
    assert torch.equal(perm.cpu(), p_ref)


# --------------------------------------------------------------- contraction family ---
NODE_CASES = [
    # (N, E, F, C, CA, CO, hasA, hasW, hasU, sorted)
    (64, 600, 128, 1, 1, 1, True, True, False, False),      # CFConv aggregate
    (64, 600, 128, 1, 1, 1, True, True, False, True),       # same, edges already grouped (perm NULL)
    (64, 600, 128, 1, 1, 1, False, True, False, False),     # segment_sum
    (40, 300, 3, 1, 1, 1, False, True, False, False),       # coordinate update, F=3 (scalar path)
    (40, 300, 1, 1, 1, 1, False, True, False, True),        # energy sum, F=1
    (50, 700, 64, 3, 1, 3, True, True, True, False),        # PaiNN direction term
    (50, 700, 64, 3, 3, 3, True, True, False, False),       # PaiNN vector term
    (50, 700, 12, 3, 3, 1, True, True, True, False),        # backward w.r.t. a 1-channel operand
    (50, 700, 20, 3, 1, 3, True, False, True, False),
    (30, 0, 16, 1, 1, 1, True, True, False, False),         # no edges -> zeros
    (2000, 60000, 128, 1, 1, 1, True, True, False, True),
    (33, 500, 6, 1, 1, 1, True, True, False, False),        # F not a multiple of 4
]


@pytest.mark.parametrize("case", NODE_CASES)
def test_contract_node_bit_exact(ops, case):
    N, E, F, C, CA, CO, hasA, hasW, hasU, srt = case
    gen = torch.Generator().manual_seed(hash(case) % 1000)
    ei = random_graph(gen, N, E, sorted_dst=srt)
    A = (torch.randn(N, F, generator=gen) if CA == 1 else torch.randn(N, CA, F, generator=gen)) if hasA else None
    W = torch.randn(E, F, generator=gen) if hasW else None
    U = torch.randn(E, C, generator=gen) if hasU else None
    topo = make_topo(ops, ei, N, sorted_side=1 if srt else None)
    dev = lambda t: None if t is None else t.to(DEV)
    got = ops["contract"]._node_raw(topo, 1, dev(A), 0, dev(W), dev(U), C, CO, F)
    want = ks.contract_node(ei[1].int(), N, A, ei[0].int(), W, U, C, CO, F)
    assert torch.equal(got.cpu().reshape(want.shape), want)
    again = ops["contract"]._node_raw(topo, 1, dev(A), 0, dev(W), dev(U), C, CO, F)
    assert torch.equal(got, again)   # run-to-run bit reproducible


@pytest.mark.parametrize("F,C,CG,CX,hasG,hasX,hasU", [
    (128, 1, 1, 1, True, True, False), (128, 1, 1, 1, True, False, False), (3, 1, 1, 1, True, False, False),
    (64, 3, 3, 1, True, True, True), (64, 3, 3, 3, True, True, False), (10, 3, 1, 1, True, True, True),
    (1, 1, 1, 1, True, False, False)])
def test_contract_edge_w_bit_exact(ops, F, C, CG, CX, hasG, hasX, hasU):
    gen = torch.Generator().manual_seed(F * 7 + C)
    N, E = 80, 900
    ei = random_graph(gen, N, E)
    mk = lambda ch: torch.randn(N, F, generator=gen) if ch == 1 else torch.randn(N, ch, F, generator=gen)
    G = mk(CG) if hasG else None
    X = mk(CX) if hasX else None
    U = torch.randn(E, C, generator=gen) if hasU else None
    topo = make_topo(ops, ei, N)
    dev = lambda t: None if t is None else t.to(DEV)
    got = ops["contract"]._edge_w_raw(topo, dev(G), 0, dev(X), 1, dev(U), C, F)
    want = ks.contract_edge_w(G, ei[0], X, ei[1], U, E, C, F)
    assert torch.equal(got.cpu(), want)


@pytest.mark.parametrize("F,C,CG,CX,hasW", [(128, 1, 1, 1, True), (64, 3, 3, 1, True), (64, 3, 3, 3, False),
                                            (12, 3, 1, 3, True), (3, 1, 1, 1, False)])
def test_contract_edge_u_close(ops, F, C, CG, CX, hasW):
    gen = torch.Generator().manual_seed(F + C)
    N, E = 80, 900
    ei = random_graph(gen, N, E)
    mk = lambda ch: torch.randn(N, F, generator=gen) if ch == 1 else torch.randn(N, ch, F, generator=gen)
    G, X = mk(CG), mk(CX)
    W = torch.randn(E, F, generator=gen) if hasW else None
    topo = make_topo(ops, ei, N)
    dev = lambda t: None if t is None else t.to(DEV)
    got = ops["contract"]._edge_u_raw(topo, dev(G), 0, dev(X), 1, dev(W), C, F)
    want = ks.contract_edge_u(G, ei[0], X, ei[1], W, E, C, F)
    # fp32 dot product of F terms in a different (fixed tree) order: |err| <= F * eps * sum|terms|
    torch.testing.assert_close(got.cpu(), want, rtol=1e-5, atol=1e-5)
    again = ops["contract"]._edge_u_raw(topo, dev(G), 0, dev(X), 1, dev(W), C, F)
    assert torch.equal(got, again)


def test_edge_gather_add_bit_exact(ops):
    gen = torch.Generator().manual_seed(3)
    N, E, F = 100, 1500, 128
    ei = random_graph(gen, N, E)
    P, Q = torch.randn(N, F, generator=gen), torch.randn(N, F, generator=gen)
    topo = make_topo(ops, ei, N)
    got = ops["contract"]._edge_gather_add_raw(topo, P.to(DEV), 0, Q.to(DEV), 1)
    assert torch.equal(got.cpu(), ks.edge_gather_add(P, ei[0], Q, ei[1]))


def test_strided_operands(ops):
    """Column slices of a wider matrix (PaiNN's split of the [E,3F] filter) are read in place."""
    gen = torch.Generator().manual_seed(5)
    N, E, F = 60, 800, 32
    ei = random_graph(gen, N, E)
    wide_w, wide_a = torch.randn(E, 3 * F, generator=gen), torch.randn(N, 3 * F, generator=gen)
    topo = make_topo(ops, ei, N)
    Wd, Ad = wide_w.to(DEV), wide_a.to(DEV)
    for k in range(3):
        got = ops["contract"]._node_raw(topo, 0, Ad[:, k * F:(k + 1) * F], 1, Wd[:, k * F:(k + 1) * F], None, 1, 1, F)
        want = ks.contract_node(ei[0].int(), N, wide_a[:, k * F:(k + 1) * F], ei[1].int(),
                                wide_w[:, k * F:(k + 1) * F], None, 1, 1, F)
        assert torch.equal(got.cpu(), want)


def test_autograd_closure_double_backward(ops):
    """d/dA, d/dW, d/dU of contract_node and their derivatives against torch autograd on CPU."""
    gen = torch.Generator().manual_seed(11)
    N, E, F = 30, 200, 16
    ei = random_graph(gen, N, E)
    A = torch.randn(N, 3, F, generator=gen)
    W = torch.randn(E, F, generator=gen)
    U = torch.randn(E, 3, generator=gen)
    probe = torch.randn(N, 3, F, generator=gen)

    def loss_fn(a, w, u, contract):
        out = contract(a, w, u)
        (ga,) = torch.autograd.grad((out * probe.to(out.device)).sum(), a, create_graph=True)
        return (ga ** 2).sum() + (out ** 2).sum()

    topo = make_topo(ops, ei, N)
    a, w, u = (t.clone().to(DEV).requires_grad_(True) for t in (A, W, U))
    loss = loss_fn(a, w, u, lambda a_, w_, u_: ops["ops"].contract_node(topo, 0, A=a_, side_a=1, W=w_, U=u_, C=3, CO=3, F=F))
    loss.backward()

    a2, w2, u2 = (t.clone().requires_grad_(True) for t in (A, W, U))

    def cpu_contract(a_, w_, u_):
        term = a_[ei[1]] * w_.unsqueeze(1) * u_.unsqueeze(-1)
        return torch.zeros(N, 3, F).index_add(0, ei[0], term)

    loss2 = loss_fn(a2, w2, u2, cpu_contract)
    loss2.backward()
    torch.testing.assert_close(loss.detach().cpu(), loss2.detach(), rtol=1e-5, atol=1e-5)
    for got, want in ((a, a2), (w, w2), (u, u2)):
        torch.testing.assert_close(got.grad.cpu(), want.grad, rtol=2e-4, atol=2e-4)


# ------------------------------------------------------------------------------ PNA ---
@pytest.mark.parametrize("N,E,F,hasR,hasM", [(40, 300, 8, False, False), (40, 300, 8, True, False),
                                              (200, 3000, 128, True, False), (10, 0, 4, False, False),
                                              (40, 300, 8, False, True), (200, 3000, 64, True, True)])
def test_pna_aggregate(ops, N, E, F, hasR, hasM):
    gen = torch.Generator().manual_seed(N + F)
    ei = random_graph(gen, N, E)
    # quantised values create exact ties for min / max
    P = torch.randint(-3, 4, (N, F), generator=gen).float() * 0.25
    Q = torch.randint(-3, 4, (N, F), generator=gen).float() * 0.25
    R = torch.randn(E, F, generator=gen) if hasR else None
    M = torch.randint(-2, 3, (E, F), generator=gen).float() * 0.5 if hasM else None   # PNAPlus multiplier (zeros too)
    topo = make_topo(ops, ei, N)
    dev = lambda t: None if t is None else t.to(DEV)
    out, cmin, cmax = ops["pna"]._pna_fwd_raw(dev(P), dev(Q), dev(R), topo, dev(M))
    want, wmin, wmax = ks.pna_aggregate_fwd(P, Q, R, ei[0], ei[1], N, M)
    F4 = out.size(1) // 4
    assert torch.equal(out.cpu()[:, F4:3 * F4], want[:, F4:3 * F4])          # min, max: exact
    assert torch.equal(cmin.cpu(), wmin) and torch.equal(cmax.cpu(), wmax)
    torch.testing.assert_close(out.cpu(), want, rtol=1e-5, atol=1e-6)        # mean, std
    gout = torch.randn(N, 4 * F, generator=gen)
    gh, gm = ops["pna"]._pna_bwd_raw(gout.to(DEV), out, cmin, cmax, dev(P), dev(Q), dev(R), topo, dev(M))
    res = ks.pna_aggregate_bwd(gout, out.cpu(), cmin.cpu(), cmax.cpu(), P, Q, R, ei[0], ei[1], N, M)
    if hasM:
        torch.testing.assert_close(gh.cpu(), res[0], rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(gm.cpu(), res[1], rtol=1e-5, atol=1e-6)
    else:
        assert gm is None
        torch.testing.assert_close(gh.cpu(), res, rtol=1e-5, atol=1e-6)


# ------------------------------------------------------------------------ attention ---
@pytest.mark.parametrize("sizes,H,D", [([37], 4, 32), ([5, 64, 1, 130, 23], 4, 32), ([300], 2, 8),
                                       ([9, 9, 9], 1, 16), ([70, 3], 4, 4), ([1024], 4, 32),
                                       # the shapes bench.py times (config 5): whole-batch and per-graph
                                       ([8192], 4, 32), ([1024] * 8, 4, 32)])
def test_attn_varlen(ops, sizes, H, D):
    gen = torch.Generator().manual_seed(sum(sizes) + H)
    N = sum(sizes)
    cu = torch.tensor([0] + list(np.cumsum(sizes)), dtype=torch.int32)
    q, k, v = (torch.randn(N, H * D, generator=gen) for _ in range(3))
    do = torch.randn(N, H * D, generator=gen)
    scale = 1.0 / np.sqrt(D)
    o, lse, *_ = ops["attention"]._attn_fwd_raw(q.to(DEV), k.to(DEV), v.to(DEV), cu.to(DEV), H, scale)
    o_ref, lse_ref = ks.attn_varlen_fwd(q, k, v, cu, H, scale)
    # fp32 online softmax vs dense softmax: north_star tolerance rel 1e-5 (+ abs floor of a few ulp)
    torch.testing.assert_close(o.cpu(), o_ref, rtol=1e-5, atol=2e-6)
    torch.testing.assert_close(lse.cpu(), lse_ref, rtol=1e-5, atol=2e-6)
    dq, dk, dv = ops["attention"]._attn_bwd_raw(do.to(DEV), o, lse, q.to(DEV), k.to(DEV), v.to(DEV),
                                               cu.to(DEV), H, scale)
    rq, rk, rv = ks.attn_varlen_bwd(do, q, k, v, cu, H, scale)
    # gradients: within 1e-5 of the dense fp64 result (relative to the tensor's scale), or as close to
    # it as the dense fp32 CPU evaluation is
    tq, tk, tv = ks.attn_varlen_bwd(do.double(), q.double(), k.double(), v.double(), cu, H, scale)
    for what, got, want, truth in (("dq", dq, rq, tq), ("dk", dk, rk, tk), ("dv", dv, rv, tv)):
        pu.assert_close_to_truth(got, want, truth, what)


def test_attn_strided_qkv(ops):
    """q,k,v as column blocks of the packed in_proj output (no copies)."""
    gen = torch.Generator().manual_seed(2)
    N, H, D = 150, 4, 32
    qkv = torch.randn(N, 3 * H * D, generator=gen)
    cu = torch.tensor([0, 50, 150], dtype=torch.int32)
    c = H * D
    qd = qkv.to(DEV)
    out = ops["ops"].attn_varlen(qd[:, :c], qd[:, c:2 * c], qd[:, 2 * c:], cu.to(DEV), H)
    ref, _ = ks.attn_varlen_fwd(qkv[:, :c], qkv[:, c:2 * c], qkv[:, 2 * c:], cu, H, 1.0 / np.sqrt(D))
    torch.testing.assert_close(out.cpu(), ref, rtol=1e-5, atol=2e-6)


def test_attn_packed_gradient_is_written_in_place(ops):
    """attn_varlen_packed(qkv): the backward kernels write dq | dk | dv straight into the column blocks of ONE
    [N, 3C] gradient (hgk_attn_varlen_bwd_ld, row stride 3C).  Output and gradient equal the three-tensor op
    bit for bit (same kernels), and the oracle within tolerance."""
    gen = torch.Generator().manual_seed(5)
    sizes, H, D = [130, 7, 300, 64], 4, 32
    N, c = sum(sizes), H * D
    cu = torch.tensor([0] + list(np.cumsum(sizes)), dtype=torch.int32)
    qkv = torch.randn(N, 3 * c, generator=gen)
    do = torch.randn(N, c, generator=gen)
    a = qkv.to(DEV).requires_grad_(True)
    out_a = ops["ops"].attn_varlen_packed(a, cu.to(DEV), H)
    out_a.backward(do.to(DEV))
    b = qkv.to(DEV).requires_grad_(True)
    out_b = ops["ops"].attn_varlen(b[:, :c], b[:, c:2 * c], b[:, 2 * c:], cu.to(DEV), H)
    out_b.backward(do.to(DEV))
    assert torch.equal(out_a, out_b)
    assert torch.equal(a.grad, b.grad)
    rq, rk, rv = ks.attn_varlen_bwd(do, qkv[:, :c], qkv[:, c:2 * c], qkv[:, 2 * c:], cu, H, 1.0 / np.sqrt(D))
    want = torch.cat([rq, rk, rv], dim=1)
    torch.testing.assert_close(a.grad.cpu(), want, rtol=1e-5, atol=1e-5 * float(want.abs().max()))


# --------------------------------------------------------------------- radius graph ---
def _canon(src, dst, S):
    order = np.lexsort((S[:, 2], S[:, 1], S[:, 0], src, dst))
    return src[order], dst[order], S[order]


@pytest.mark.parametrize("case", ["h2_loop0", "h2_loop1", "cr_loop0", "cr_loop1", "tri_loop0", "tri_loop1",
                                  "h2_big_loop0", "h2_big_loop1"])
def test_radius_graph_pbc_vs_reference_golden(ops, golden, case):
    """Edge lists of the REAL reference RadiusGraphPBC (tests/golden), bit-exact after canonical sort."""
    c = golden["pbc"][case]
    n = c["pos"].size(0)
    g = ops["ops"].radius_graph(c["pos"].to(DEV), torch.tensor([0, n]), c["r"], c["cap"], c["loop"],
                                cell=c["cell"].reshape(1, 3, 3), pbc=c["pbc"].reshape(1, 3))
    src, dst, S = _canon(g["src"].cpu().numpy(), g["dst"].cpu().numpy(), g["shift_int"].cpu().numpy())
    ref_shift = c["edge_shifts"].numpy()
    ref_S = np.rint(ref_shift @ np.linalg.inv(c["cell"].numpy().astype(np.float64))).astype(np.int64)
    rs, rd, rS = _canon(c["edge_index"][0].numpy(), c["edge_index"][1].numpy(), ref_S)
    assert np.array_equal(src, rs) and np.array_equal(dst, rd) and np.array_equal(S, rS)
    assert torch.allclose(g["edge_shifts"].cpu(), g["shift_int"].cpu().float() @ c["cell"], atol=1e-5)


def _random_cells(gen, B):
    cells = []
    for _ in range(B):
        a = 6.0 + 3.0 * torch.rand(3, generator=gen)
        cell = torch.diag(a)
        cell[1, 0] = 1.5 * torch.rand(1, generator=gen).item()
        cell[2, 0] = 0.8 * torch.rand(1, generator=gen).item()
        cell[2, 1] = 1.1 * torch.rand(1, generator=gen).item()
        cells.append(cell)
    return torch.stack(cells)


@pytest.mark.parametrize("cap,loop,pbcs", [(0, False, (1, 1, 1)), (12, False, (1, 1, 1)), (0, True, (1, 1, 0)),
                                           (20, True, (1, 0, 0)), (0, False, (0, 0, 0))])
def test_radius_graph_ase_batched_bit_exact(ops, cap, loop, pbcs):
    gen = torch.Generator().manual_seed(17 + cap)
    B = 5
    sizes = [int(s) for s in torch.randint(1, 60, (B,), generator=gen)]
    ptr = torch.tensor([0] + list(np.cumsum(sizes)))
    cells = _random_cells(gen, B)
    pos = torch.cat([(torch.rand(n, 3, generator=gen) * 1.6 - 0.3) @ cells[b] for b, n in enumerate(sizes)])
    pbc = torch.tensor([pbcs] * B, dtype=torch.bool)
    g = ops["ops"].radius_graph(pos.to(DEV), ptr, 3.1, cap, loop, cell=cells, pbc=pbc)
    src, dst, S, dist, rowptr = ks.radius_graph(pos, ptr, 3.1, cap, loop, cells, pbc, mode=0)
    assert torch.equal(g["rowptr"].cpu(), rowptr)
    assert torch.equal(g["src"].cpu(), src) and torch.equal(g["dst"].cpu(), dst)
    assert torch.equal(g["shift_int"].cpu(), S)
    assert torch.equal(g["dist"].cpu(), dist)     # fp64 distances, same operation order


@pytest.mark.parametrize("cap", [0, 8, 32])
def test_radius_graph_cluster_bit_exact(ops, cap):
    gen = torch.Generator().manual_seed(23 + cap)
    sizes = [18, 3, 29, 1, 22, 40]
    ptr = torch.tensor([0] + list(np.cumsum(sizes)))
    pos = torch.cat([torch.rand(n, 3, generator=gen) * 6.0 for n in sizes])
    g = ops["ops"].radius_graph(pos.to(DEV), ptr, 3.0, cap, False)
    src, dst, S, dist, rowptr = ks.radius_graph(pos, ptr, 3.0, cap, False, mode=1)
    assert torch.equal(g["src"].cpu(), src) and torch.equal(g["dst"].cpu(), dst)
    assert torch.equal(g["rowptr"].cpu(), rowptr)


def test_radius_graph_fesi_full_size_properties(ops):
    """BASELINE config 5 at full size (8 x 1024-atom BCC cells, r=5): size-independent properties."""
    gen = torch.Generator().manual_seed(5)
    a, n = 2.8, 8
    grid = torch.stack(torch.meshgrid(*[torch.arange(n)] * 3, indexing="ij"), -1).reshape(-1, 3).float()
    base = torch.cat([grid, grid + 0.5]) * a
    B = 8
    pos = torch.cat([base + 0.05 * torch.randn(base.shape, generator=gen) for _ in range(B)])
    ptr = torch.arange(B + 1) * base.size(0)
    cell = (torch.eye(3) * a * n).repeat(B, 1, 1)
    pbc = torch.ones(B, 3, dtype=torch.bool)
    g = ops["ops"].radius_graph(pos.to(DEV), ptr, 5.0, 64, False, cell=cell, pbc=pbc)
    src, dst = g["src"].long(), g["dst"].long()
    deg = torch.bincount(dst, minlength=pos.size(0))
    # BCC shells 8+6+12+24+8 = 58 inside 5 A; the 0.05 A jitter pushes a few 5th-shell pairs (4.85 A) out
    assert int(deg.max()) <= 58 and int(deg.min()) >= 48 and float(deg.float().mean()) > 56.5
    # completeness: brute force over the 27 images in fp64 for the first crystal
    p64 = pos[:1024].double().to(DEV)
    shifts = torch.stack(torch.meshgrid(*[torch.arange(-1, 2)] * 3, indexing="ij"), -1).reshape(-1, 3).double().to(DEV) * (a * n)
    D = (p64[None, :, None, :] - p64[:, None, None, :]) + shifts[None, None, :, :]     # [i, j, S, 3]
    dist = torch.sqrt(D[..., 0] * D[..., 0] + D[..., 1] * D[..., 1] + D[..., 2] * D[..., 2])
    hit = dist < 5.0
    hit[torch.arange(1024), torch.arange(1024), 13] = False                            # i == j, S == 0
    assert int(hit.sum()) == int((dst < 1024).sum())
    assert torch.equal(hit.sum(dim=(0, 2)).to(torch.int64), deg[:1024])
    assert torch.equal(dst, torch.sort(dst).values)                # grouped by destination
    assert bool(((src // 1024) == (dst // 1024)).all())            # never crosses graphs
    vec = pos.to(DEV)[dst] - pos.to(DEV)[src] + g["edge_shifts"]
    d = vec.norm(dim=1)
    assert float(d.max()) < 5.0 + 1e-4
    # symmetry: (i, j, S) present  <=>  (j, i, -S) present
    key = lambda s, t, S: ((s * 8192 + t) * 27 + (S[:, 0] + 1) * 9 + (S[:, 1] + 1) * 3 + (S[:, 2] + 1))
    S = g["shift_int"].long()
    assert torch.equal(torch.sort(key(src, dst, S)).values, torch.sort(key(dst, src, -S)).values)


# ------------------------------------------------------------- fused SchNet filter ---
@pytest.mark.parametrize("E,R", [(1, 8), (127, 8), (1000, 0), (40000, 8), (474078, 8)])   # last: config 5's E
def test_schnet_filter_fwd(ops, E, R):
    """tcgen05 3xTF32 kernel vs the reference op sequence in fp32 on the CPU (and fp64 as the truth):
    the kernel must be as close to fp64 as the fp32 CPU path is, within rel 1e-5 of it."""
    from hydragnn.ops import filter as filt

    gen = torch.Generator().manual_seed(E + R)
    F, G, cutoff = 128, 50, 5.0
    dist = torch.rand(E, generator=gen) * 4.9 + 0.05
    raw = torch.rand(E, R, generator=gen) * 0.2 if R else None
    offset = torch.linspace(0.0, cutoff, G)
    coeff = -0.5 / float(offset[1] - offset[0]) ** 2
    w0g = torch.randn(F, G, generator=gen) / G ** 0.5
    w0r = torch.randn(F, R, generator=gen) if R else None
    b0, b1 = torch.randn(F, generator=gen) * 0.1, torch.randn(F, generator=gen) * 0.1
    w1 = torch.randn(F, F, generator=gen) / F ** 0.5
    dev = lambda t: None if t is None else t.to(DEV)
    w0p = torch.zeros(F, 64)
    w0p[:, :G] = w0g
    if R:
        w0p[:, 52:52 + R] = w0r
    w, h = filt._filter_fwd_raw(dev(dist), dev(raw), dev(offset), coeff, cutoff, dev(w0p), dev(b0), dev(w1), dev(b1))
    w_ref, h_ref = ks.schnet_filter_fwd(dist, raw, offset, coeff, cutoff, w0g, w0r, b0, w1, b1)
    torch.testing.assert_close(h.cpu(), h_ref, rtol=1e-5, atol=2e-6)
    torch.testing.assert_close(w.cpu(), w_ref, rtol=1e-5, atol=2e-6)
    dd = lambda t: None if t is None else t.double()
    w64, _ = ks.schnet_filter_fwd(dd(dist), dd(raw), dd(offset), coeff, cutoff, dd(w0g), dd(w0r), dd(b0), dd(w1), dd(b1))
    err_kernel = (w.cpu().double() - w64).abs().max().item()
    err_cpu32 = (w_ref.double() - w64).abs().max().item()
    assert err_kernel < 4 * err_cpu32 + 1e-6, (err_kernel, err_cpu32)


@pytest.mark.parametrize("E,R", [(70, 8), (5000, 0), (40000, 8), (474078, 8)])
def test_schnet_filter_backward_kernels(ops, E, R):
    """hgk_schnet_filter_bwd + hgk_wgrad_tn (through the autograd Function) vs torch autograd of the
    reference op sequence on the CPU."""
    from hydragnn.ops import filter as filt

    gen = torch.Generator().manual_seed(E * 3 + R)
    F, G, cutoff = 128, 50, 5.0
    dist = torch.rand(E, generator=gen) * 4.9 + 0.05
    raw = torch.rand(E, R, generator=gen) * 0.2 if R else None
    offset = torch.linspace(0.0, cutoff, G)
    coeff = -0.5 / float(offset[1] - offset[0]) ** 2
    params = [torch.randn(F, G, generator=gen) / G ** 0.5, torch.randn(F, max(R, 1), generator=gen),
              torch.randn(F, generator=gen) * 0.1, torch.randn(F, F, generator=gen) / F ** 0.5,
              torch.randn(F, generator=gen) * 0.1]
    gout = torch.randn(E, F, generator=gen) / E ** 0.5

    def cpu_grads(dtype):
        cast = lambda t: None if t is None else t.to(dtype)
        leaves = [t.clone().to(dtype).requires_grad_(True) for t in params]
        w_ref, _ = ks.schnet_filter_fwd(cast(dist), cast(raw), cast(offset), coeff, cutoff, leaves[0],
                                        leaves[1][:, :R] if R else None, leaves[2], leaves[3], leaves[4])
        w_ref.backward(cast(gout))
        return [t.grad for t in leaves]

    ref32, truth = cpu_grads(torch.float32), cpu_grads(torch.float64)
    dev = lambda t: None if t is None else t.to(DEV)
    gpu = [t.clone().to(DEV).requires_grad_(True) for t in params]
    w = filt.schnet_filter(dev(dist), dev(raw), dev(offset), coeff, cutoff, gpu[0], gpu[1][:, :R] if R else None,
                           gpu[2], gpu[3], gpu[4])
    w.backward(dev(gout))
    for k, g in enumerate(gpu):
        if k == 1 and not R:
            continue
        pu.assert_close_to_truth(g.grad, ref32[k], truth[k], f"param {k}")


# ----------------------------------------------------------------------- edge cases ---
def test_attn_empty_and_tiny_segments(ops):
    """cu_seqlens with empty segments, single-node segments and a segment just over one tile."""
    gen = torch.Generator().manual_seed(9)
    sizes = [0, 1, 0, 129, 2, 0]
    N, H, D = sum(sizes), 4, 32
    cu = torch.tensor([0] + list(np.cumsum(sizes)), dtype=torch.int32)
    q, k, v, do = (torch.randn(N, H * D, generator=gen) for _ in range(4))
    scale = D ** -0.5
    o, lse, *_ = ops["attention"]._attn_fwd_raw(q.to(DEV), k.to(DEV), v.to(DEV), cu.to(DEV), H, scale)
    o_ref, lse_ref = ks.attn_varlen_fwd(q, k, v, cu, H, scale)
    torch.testing.assert_close(o.cpu(), o_ref, rtol=1e-5, atol=2e-6)
    torch.testing.assert_close(lse.cpu(), lse_ref, rtol=1e-5, atol=2e-6)
    dq, dk, dv = ops["attention"]._attn_bwd_raw(do.to(DEV), o, lse, q.to(DEV), k.to(DEV), v.to(DEV), cu.to(DEV), H, scale)
    rq, rk, rv = ks.attn_varlen_bwd(do, q, k, v, cu, H, scale)
    tq, tk, tv = ks.attn_varlen_bwd(do.double(), q.double(), k.double(), v.double(), cu, H, scale)
    for what, got, want, truth in (("dq", dq, rq, tq), ("dk", dk, rk, tk), ("dv", dv, rv, tv)):
        pu.assert_close_to_truth(got, want, truth, what)


def test_radius_graph_no_edges_and_single_atoms(ops):
    """Graphs whose atoms are all farther apart than the cutoff, and one-atom graphs: E = 0."""
    pos = torch.tensor([[0.0, 0, 0], [10.0, 0, 0], [0, 10.0, 0], [50.0, 50, 50]])
    ptr = torch.tensor([0, 3, 4])
    g = ops["ops"].radius_graph(pos.to(DEV), ptr, 2.0, 8, False)
    assert g["edge_index"].shape == (2, 0) and int(g["rowptr"][-1]) == 0
    cell = (torch.eye(3) * 30.0).repeat(2, 1, 1)
    g = ops["ops"].radius_graph(pos.to(DEV), ptr, 2.0, 8, False, cell=cell, pbc=torch.ones(2, 3, dtype=torch.bool))
    assert g["edge_index"].shape == (2, 0)
    # one atom in a small periodic cell sees its own images (ASE semantics), never itself at S = 0
    g = ops["ops"].radius_graph(torch.zeros(1, 3, device=DEV), torch.tensor([0, 1]), 1.5, 0, False,
                                cell=torch.eye(3).reshape(1, 3, 3), pbc=torch.ones(1, 3, dtype=torch.bool))
    src, dst, S, dist, _ = ks.radius_graph(torch.zeros(1, 3), torch.tensor([0, 1]), 1.5, 0, False,
                                           torch.eye(3).reshape(1, 3, 3), torch.ones(1, 3, dtype=torch.bool), mode=0)
    assert torch.equal(g["shift_int"].cpu(), S) and g["src"].numel() == 18      # 6 faces + 12 edges of the cube


def test_segment_reductions_with_isolated_rows_and_long_segments(ops):
    """Rows without edges give exact zeros; one hub row with thousands of edges keeps CPU order."""
    gen = torch.Generator().manual_seed(4)
    N, F = 300, 128
    dst = torch.cat([torch.full((5000,), 7), torch.randint(100, 200, (3000,), generator=gen)])
    src = torch.randint(0, N, (dst.numel(),), generator=gen)
    ei = torch.stack([src, dst])
    x, W = torch.randn(N, F, generator=gen), torch.randn(dst.numel(), F, generator=gen)
    topo = make_topo(ops, ei, N)
    got = ops["contract"]._node_raw(topo, 1, x.to(DEV), 0, W.to(DEV), None, 1, 1, F).cpu()
    want = ks.contract_node(ei[1].int(), N, x, ei[0].int(), W, None, 1, 1, F)
    assert torch.equal(got, want)
    assert float(got[:7].abs().max()) == 0.0 and float(got[200:].abs().max()) == 0.0


def test_attn_varlen_double_backward(ops):
    """create_graph=True through the attention op (a force loss on a GPS model whose node features
    depend on the positions): the backward switches to differentiable torch ops; first and second
    derivatives agree with a dense torch attention."""
    from hydragnn.ops import attn_varlen

    gen = torch.Generator().manual_seed(21)
    sizes, H, D = [37, 5, 64], 4, 32
    N = sum(sizes)
    cu = torch.tensor([0] + list(torch.tensor(sizes).cumsum(0)), dtype=torch.int32, device=DEV)
    base = [torch.randn(N, H * D, generator=gen).to(DEV) for _ in range(3)]

    def dense(q, k, v):
        outs, b = [], 0
        for L in sizes:
            qs, ks, vs = (t[b:b + L].reshape(L, H, D).transpose(0, 1) for t in (q, k, v))
            p = torch.softmax((qs @ ks.transpose(1, 2)) * D ** -0.5, dim=-1)
            outs.append((p @ vs).transpose(0, 1).reshape(L, H * D))
            b += L
        return torch.cat(outs)

    res = []
    for fn in (lambda q, k, v: attn_varlen(q, k, v, cu, H), dense):
        q, k, v = (t.clone().requires_grad_(True) for t in base)
        e = (fn(q, k, v) ** 2).sum()
        gq, gk = torch.autograd.grad(e, (q, k), create_graph=True)
        loss = (gq * gq).sum() + (gk * v).sum()
        loss.backward()
        res.append((e.detach(), gq.detach(), loss.detach(), q.grad, k.grad, v.grad))
    for a, b in zip(*res):
        torch.testing.assert_close(a, b, rtol=2e-4, atol=2e-4)


@pytest.mark.parametrize("n,c,strided", [(8192, 128, False), (5000, 384, False), (1, 7, False), (333, 256, True),
                                         (0, 16, False)])
def test_colsum(ops, n, c, strided):
    """hgk_colsum (bias gradient of a node-row Linear): fp64 reference, run-to-run identical bits."""
    from hydragnn.ops import colsum

    gen = torch.Generator().manual_seed(n + c)
    full = torch.randn(n, c * (2 if strided else 1), generator=gen).to(DEV)
    x = full[:, :c]
    got = colsum(x)
    want = x.double().sum(0)
    assert got.shape == (c,)
    torch.testing.assert_close(got.double(), want, rtol=1e-5, atol=1e-4 * max(1.0, n ** 0.5))
    assert torch.equal(got, colsum(x))


@pytest.mark.parametrize("N,E", [(50, 700), (8192, 474078)])
def test_cfconv_gathered_backward_kernels(ops, N, E):
    """hgk_cfconv_filter_bwd / hgk_cfconv_wgrad (filter gradient gW[e] = g[dst(e)] * x1[src(e)] formed
    inside the kernel) against the two-step path they replace: hgk_contract_edge_w writes gW, then
    hgk_schnet_filter_bwd / hgk_wgrad_tn read it.  Same arithmetic on the same values: bit-exact."""
    from hydragnn.ops import filter as filt

    gen = torch.Generator().manual_seed(N + E)
    F, cutoff = 128, 5.0
    ei = random_graph(gen, N, E, sorted_dst=True)
    topo = make_topo(ops, ei, N, sorted_side=1)
    g = torch.randn(N, F, generator=gen).to(DEV)
    x1 = torch.randn(N, F, generator=gen).to(DEV)
    h = (torch.rand(E, F, generator=gen) * 3.0 - 0.6).to(DEV)        # the range of ssp(z) = softplus(z) - ln 2
    dist = (torch.rand(E, generator=gen) * 4.9 + 0.05).to(DEV)
    w1t = (torch.randn(F, F, generator=gen) / F ** 0.5).to(DEV)
    gw = ops["contract"]._edge_w_raw(topo, g, 1, x1, 0, None, 1, F)
    dz0_ref = filt._filter_bwd_raw(gw, h, dist, cutoff, w1t)
    dz0 = filt._cfconv_filter_bwd_raw(g, x1, topo, h, dist, cutoff, w1t)
    assert torch.equal(dz0, dz0_ref)
    gw1_ref, gb1_ref = filt._wgrad_raw(gw, dist, cutoff, b=h)
    gw1, gb1 = filt._cfconv_wgrad_raw(g, x1, topo, dist, cutoff, h)
    assert torch.equal(gw1, gw1_ref) and torch.equal(gb1, gb1_ref)
    # and against plain torch on the CPU
    gw_cpu = g.cpu()[ei[1]] * x1.cpu()[ei[0]]
    ref = ks.filter_bwd(gw_cpu, h.cpu(), dist.cpu(), cutoff, w1t.cpu())
    truth = ks.filter_bwd(gw_cpu.double(), h.cpu().double(), dist.cpu().double(), cutoff, w1t.cpu().double())
    pu.assert_close_to_truth(dz0, ref, truth, "dz0")


# -------------------------------------------------------------------- edge geometry ---
@pytest.mark.parametrize("normalize,eps,with_shifts", [(False, 1e-9, False), (True, 1e-9, True), (True, 1.0, False)])
def test_edge_geometry_kernels(ops, normalize, eps, with_shifts):
    """hgk_edge_geom_fwd / _bwd / _bwd2 vs their restatement, and the operator (value, first derivative and
    the second-order term of the force loss) vs torch autograd through the reference's op chain in fp64."""
    import sys, os
    sys.path.insert(0, os.path.dirname(__file__))
    from test_host_logic_cpu import _geometry_case
    from hydragnn.ops import geometry

    gen = torch.Generator().manual_seed(11)
    N, E = 300, 5000
    pos = torch.randn(N, 3, generator=gen) * 2
    src = torch.randint(0, N, (E,), generator=gen, dtype=torch.int32)
    dst = (src + 1 + torch.randint(0, N - 1, (E,), generator=gen, dtype=torch.int32)) % N      # no self pairs
    shifts = torch.randn(E, 3, generator=gen) * 0.1 if with_shifts else None
    d = lambda t: None if t is None else t.to(DEV)
    vec, length, unit = geometry._fwd_raw(d(pos), d(src), d(dst), d(shifts), eps, normalize)
    rv, rl, ru = ks.edge_geom_fwd(pos, src, dst, shifts, eps, normalize)
    torch.testing.assert_close(vec.cpu(), rv, rtol=0, atol=0)             # sub / add with the reference's roundings
    torch.testing.assert_close(length.cpu(), rl, rtol=1e-6, atol=1e-7)
    if normalize:
        torch.testing.assert_close(unit.cpu(), ru, rtol=1e-6, atol=1e-7)
    g_vec, g_unit, h = (torch.randn(E, 3, generator=gen) for _ in range(3))
    g_len = torch.randn(E, generator=gen)
    gu = g_unit if normalize else None
    t = geometry._bwd_raw(d(g_vec), d(g_len), d(gu), vec, length, eps)
    torch.testing.assert_close(t.cpu(), ks.edge_geom_bwd(g_vec, g_len, gu, rv, rl, eps), rtol=1e-5, atol=1e-5)
    got = geometry._bwd2_raw(d(h), d(g_len), d(gu), vec, length, eps)
    want = ks.edge_geom_bwd2(h.double(), g_len.double(), None if gu is None else gu.double(), rv.double(), rl.double(), eps)
    for a, b, what in zip(got, want, ("d g_len", "d g_unit", "d vec")):
        scale = float(b.abs().max())
        torch.testing.assert_close(a.cpu().double(), b, rtol=1e-4, atol=1e-5 * scale, msg=lambda m, what=what: f"{what}: {m}")
    # the operator end to end, second order included
    truth = _geometry_case(torch.float64, "cpu", normalize, eps, with_shifts, use_op=False)
    gpu = _geometry_case(torch.float32, DEV, normalize, eps, with_shifts, use_op=True)
    for a, b, what in zip(gpu, truth, ("value", "first derivative", "second-order term")):
        scale = float(b.abs().max())
        torch.testing.assert_close(a.cpu().double(), b, rtol=1e-4, atol=1e-5 * scale, msg=lambda m, what=what: f"{what}: {m}")
