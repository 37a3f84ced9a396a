"""Host-side logic of the product stacks (weight splitting, CSR sides, the autograd closure used
by the force loss) against the golden outputs of the REAL reference modules -- on CPU, with the
C-ABI launchers swapped for oracle/kernel_spec.py (tests/cpu_patch.py).  The kernels themselves
are checked by the `-m gpu` tests."""
import pytest
import torch

import parity_util as pu
from cpu_patch import cpu_kernels

CASES = ["schnet_plain", "schnet_equiv", "schnet_gps", "egnn_gps_edge", "egnn_equiv", "painn_forces",
         "schnet_forces", "egnn_forces", "pna_multihead", "pna_edge",
         # GPS in train mode (BatchNorm batch statistics), dropout = 0: tests/golden/make_golden_train.py
         "schnet_gps_train", "egnn_gps_edge_train", "painn_gps_train", "schnet_gps_train_wide",
         "egnn_gps_train_wide"]


@pytest.mark.parametrize("name", CASES)
def test_product_stack_matches_reference_on_cpu(name):
    with cpu_kernels():
        res = pu.run_golden_case(name, device="cpu")
    pu.assert_case_close(res, truth=pu.oracle_truth_case(name))


def test_batched_laplacian_pe_matches_per_sample():
    """preprocess.load_data: one batched eigendecomposition per graph size == the per-sample routine."""
    import types

    from hydragnn.preprocess import load_data as ld

    gen = torch.Generator().manual_seed(0)
    samples = []
    for n in (5, 5, 7, 1, 5, 7, 12):
        ei = torch.randint(0, n, (2, 3 * n), generator=gen) if n > 1 else torch.zeros(2, 0, dtype=torch.long)
        samples.append(types.SimpleNamespace(num_nodes=n, edge_index=ei))
    ld._laplacian_pe_batched(samples, 3, torch.device("cpu"))
    for d in samples:
        torch.testing.assert_close(d.pe, ld._laplacian_pe(d.edge_index, d.num_nodes, 3), rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("n,k", [(12, 4), (40, 8), (130, 8)])
def test_laplacian_pe_matches_pyg_transform_up_to_sign(n, k):
    """preprocess.load_data._laplacian_pe against the restatement of PyG's AddLaplacianEigenvectorPE
    (oracle/pyg_lite.py; reference: serialized_dataset_loader.py:90-94,183-189) on undirected radius
    graphs of random points (generic spectrum: no repeated eigenvalues).  PyG multiplies every
    eigenvector by a random sign, the product fixes the sign (first non-zero entry positive), so the
    columns agree up to sign and rel_pe = |pe_i - pe_j|, which the convs consume, agrees exactly."""
    from hydragnn.preprocess import load_data as ld
    from oracle import pyg_lite as pyg

    for seed in range(n, n + 50):       # a connected graph: a repeated eigenvalue 0 has no unique eigenvectors
        gen = torch.Generator().manual_seed(seed)
        pos = torch.rand(n, 3, generator=gen) * (n / 0.4) ** (1.0 / 3.0)
        ei = pyg.radius_graph_cluster(pos, 2.2, None, False, 32)
        ei = torch.unique(torch.cat([ei, ei.flip(0)], dim=1), dim=1)      # undirected, no duplicates
        A = torch.zeros(n, n, dtype=torch.float64)
        A[ei[0], ei[1]] = 1.0
        lap = torch.diag(A.sum(1)) - A
        vals = torch.linalg.eigvalsh(lap)
        if float(vals[1]) > 1e-6 and float((vals[1:k + 2] - vals[:k + 1]).min()) > 1e-4:
            break
    else:
        pytest.skip("no connected sample found")
    ours = ld._laplacian_pe(ei, n, k)
    ref = pyg.laplacian_eigenvector_pe(ei, n, k)
    sign = torch.sign((ours * ref).sum(0))
    assert bool((sign != 0).all())
    torch.testing.assert_close(ours, ref * sign, rtol=1e-4, atol=2e-5)
    rel_ours = (ours[ei[0]] - ours[ei[1]]).abs()
    rel_ref = (ref[ei[0]] - ref[ei[1]]).abs()
    torch.testing.assert_close(rel_ours, rel_ref, rtol=1e-4, atol=2e-5)
    # the fixed sign convention: first entry with |v| > 1e-9 is positive
    for j in range(k):
        col = ours[:, j]
        assert float(col[torch.nonzero(col.abs() > 1e-9)[0, 0]]) > 0


def test_attention_double_backward_in_row_blocks(monkeypatch):
    """ops.attention under create_graph=True (force loss through a GPS layer): the differentiable backward
    works in blocks of query rows under checkpointing; first and second derivatives equal those of the
    dense torch attention, for a whole-batch segment longer than one block and for ragged segments."""
    from hydragnn.ops import attention
    from oracle import kernel_spec as ks

    monkeypatch.setattr(attention, "_DENSE_BLOCK", 8)
    gen = torch.Generator().manual_seed(4)
    H, D = 2, 4
    for sizes in ([29], [5, 0, 13, 9]):
        N = sum(sizes)
        cu = torch.tensor([0] + list(torch.tensor(sizes).cumsum(0)), dtype=torch.int32)
        base = [torch.randn(N, H * D, generator=gen, dtype=torch.float64) for _ in range(3)]
        w = torch.randn(N, H * D, generator=gen, dtype=torch.float64)

        def second_order(fn):
            q, k, v = (t.clone().requires_grad_(True) for t in base)
            o = fn(q, k, v)
            gq, gk, gv = torch.autograd.grad((o * w).sum(), (q, k, v), create_graph=True)
            loss = (gq ** 2).sum() + (gk * gv).sum() + o.pow(3).sum()
            return [o.detach(), gq.detach()] + list(torch.autograd.grad(loss, (q, k, v)))

        with cpu_kernels():
            got = second_order(lambda q, k, v: attention.attn_varlen(q, k, v, cu, H))
        want = second_order(lambda q, k, v: ks.attn_varlen_fwd_autograd(q, k, v, cu, H, D ** -0.5)[0])
        for a, b in zip(got, want):
            torch.testing.assert_close(a, b, rtol=1e-9, atol=1e-11)


def _geometry_case(dtype, device, normalize, eps, with_shifts, use_op):
    """value, d/dpos and d/dpos of |d/dpos|^2 (second order) of a scalar function of the edge geometry."""
    from hydragnn.graph import topo_of
    from hydragnn.utils.model import get_edge_vectors_and_lengths
    from oracle import hydragnn_ref as ora

    gen = torch.Generator().manual_seed(3)
    N, E = 9, 40
    pos = (torch.randn(N, 3, generator=gen, dtype=torch.float64) * 1.5).to(dtype).to(device).requires_grad_(True)
    ei = torch.stack([torch.randint(0, N, (E,), generator=gen), torch.randint(0, N, (E,), generator=gen)])
    ei = ei[:, ei[0] != ei[1]].to(device)
    shifts = (torch.randn(ei.size(1), 3, generator=gen, dtype=torch.float64) * 0.1).to(dtype).to(device) if with_shifts else None
    w = torch.randn(ei.size(1), 3, generator=gen, dtype=torch.float64).to(dtype).to(device)
    if use_op:
        vec, length = get_edge_vectors_and_lengths(pos, ei, shifts, normalize=normalize, eps=eps)
    else:           # the reference always adds a shift tensor (zeros without PBC)
        vec, length = ora.get_edge_vectors_and_lengths(pos, ei, torch.zeros_like(w) if shifts is None else shifts,
                                                       normalize=normalize, eps=eps)
    value = (torch.sin(vec) * w).sum() + (length ** 2 * w[:, :1]).sum() + (vec * w).sum() ** 2 * 0.01
    (g,) = torch.autograd.grad(value, pos, create_graph=True)
    (gg,) = torch.autograd.grad((g ** 2).sum(), pos)
    return value.detach(), g.detach(), gg.detach()


@pytest.mark.parametrize("normalize,eps,with_shifts", [(False, 1e-9, False), (True, 1e-9, True), (True, 1.0, False)])
def test_edge_geometry_operator_first_and_second_derivatives(normalize, eps, with_shifts):
    """ops.geometry (forward, vector-Jacobian product and its hand-written derivative, over kernel_spec on the
    CPU) against torch autograd through the reference's op chain, in fp64: value, forces-like first
    derivative and the second-order term the force loss differentiates."""
    from hydragnn.ops import geometry

    want = _geometry_case(torch.float64, "cpu", normalize, eps, with_shifts, use_op=False)
    with cpu_kernels():
        got = _geometry_case(torch.float32, "cpu", normalize, eps, with_shifts, use_op=True)
    # the operator runs in fp32 (kernel_spec's segmented sums are fp32): a wrong term in a derivative formula
    # would show up at order one, rounding at 1e-6
    for a, b, what in zip(got, want, ("value", "first derivative", "second-order term")):
        scale = float(b.abs().max())
        torch.testing.assert_close(a.double(), b, rtol=1e-4, atol=1e-5 * scale, msg=lambda m, what=what: f"{what}: {m}")
