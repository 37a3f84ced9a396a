"""RadiusGraphPBC's cutoff expansion and artificial-edge fallback
(reference preprocess/graph_samples_checks_and_updates.py:163-221, 284-307) against fixtures made
by the reference's own class (tests/golden/make_golden_pbc_retry.py):

  * CPU: the oracle restatement, and the product's batched host logic (`ops.radius_graph_pbc`) with
    the kernels swapped for the oracle (tests/cpu_patch.py);
  * GPU: the product end to end through the C ABI, one sample at a time (exact edge lists) and all
    samples in one batched call (against the oracle sample by sample, same numpy seed).
"""
import os

import numpy as np
import pytest
import torch

from oracle import hydragnn_ref as ora

HERE = os.path.dirname(os.path.abspath(__file__))
FX = torch.load(os.path.join(HERE, "golden", "pbc_retry_fixtures.pt"), weights_only=False)
CASES = sorted(FX)
DEV = "cuda"


def _check_single(c, g):
    """Edge list of one sample, in order (the fixtures have no distance ties)."""
    assert torch.equal(g["edge_index"].cpu(), c["edge_index"])
    torch.testing.assert_close(g["edge_shifts"].cpu(), c["edge_shifts"], rtol=0, atol=1e-5)
    assert g["edge_ptr"].tolist() == [0, c["edge_index"].size(1)]
    assert (g["rowptr"] is None) == bool(c["artificial"])
    assert float(g["cutoff"][0]) == pytest.approx(c["r"] * 1.25 ** c["expansions"])


@pytest.mark.parametrize("case", CASES)
def test_oracle_retry_matches_reference(case):
    c = FX[case]
    np.random.seed(c["seed"])
    ei, S, d, cutoff = ora.radius_graph_pbc_full(c["pos"].numpy(), c["cell"].numpy(), c["pbc"].numpy(),
                                                 c["r"], c["cap"], c["loop"])
    assert np.array_equal(ei, c["edge_index"].numpy())
    assert np.allclose(torch.from_numpy(S).float() @ c["cell"], c["edge_shifts"], atol=1e-5)
    assert cutoff == pytest.approx(c["r"] * 1.25 ** c["expansions"])
    if c["artificial"]:
        k = ei.shape[1] - int((d < cutoff - 2e-8).sum())
        assert k >= 1 and np.allclose(d[-k:], cutoff - 1e-8)


@pytest.mark.parametrize("case", CASES)
def test_host_logic_retry_on_cpu(case):
    from cpu_patch import cpu_kernels
    from hydragnn.ops import radius as radius_mod

    c = FX[case]
    n = c["pos"].size(0)
    np.random.seed(c["seed"])
    with cpu_kernels():
        g = radius_mod.radius_graph_pbc(c["pos"], torch.tensor([0, n]), c["r"], c["cap"], c["loop"],
                                        cell=c["cell"][None], pbc=c["pbc"][None])
    _check_single(c, g)


def _batched(cases, loop, r, cap):
    pos = torch.cat([FX[k]["pos"] for k in cases])
    ptr = torch.tensor([0] + list(np.cumsum([FX[k]["pos"].size(0) for k in cases])))
    cell = torch.stack([FX[k]["cell"] for k in cases])
    pbc = torch.stack([FX[k]["pbc"] for k in cases])
    np.random.seed(99)
    want = []
    for k in cases:
        c = FX[k]
        ei, S, d, cutoff = ora.radius_graph_pbc_full(c["pos"].numpy(), c["cell"].numpy(), c["pbc"].numpy(),
                                                     r, cap, loop)
        want.append((torch.from_numpy(ei), torch.from_numpy(S).float() @ c["cell"], cutoff))
    return pos, ptr, cell, pbc, want


def _check_batched(g, ptr, want):
    ep = g["edge_ptr"].tolist()
    for k, (ei, sh, cutoff) in enumerate(want):
        a, b = ep[k], ep[k + 1]
        assert torch.equal(g["edge_index"][:, a:b].cpu() - int(ptr[k]), ei), k
        torch.testing.assert_close(g["edge_shifts"][a:b].cpu(), sh, rtol=0, atol=1e-5)
        assert float(g["cutoff"][k]) == pytest.approx(cutoff)
    assert ep[-1] == g["edge_index"].size(1)


BATCH = [k for k in CASES if k.endswith("loop0")]


@pytest.mark.parametrize("r,cap", [(2.0, 100000), (1.5, 3), (2.5, 100000)])
def test_host_logic_retry_batched_on_cpu(r, cap):
    """Samples that need 0, 1, 2 expansions and artificial edges in ONE batched call."""
    from cpu_patch import cpu_kernels
    from hydragnn.ops import radius as radius_mod

    pos, ptr, cell, pbc, want = _batched(BATCH, False, r, cap)
    np.random.seed(99)
    with cpu_kernels():
        g = radius_mod.radius_graph_pbc(pos, ptr, r, cap, False, cell=cell, pbc=pbc)
    _check_batched(g, ptr, want)


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_product_retry_matches_reference(case, libhgk):
    from hydragnn.ops import radius_graph_pbc

    c = FX[case]
    n = c["pos"].size(0)
    np.random.seed(c["seed"])
    g = radius_graph_pbc(c["pos"].to(DEV), torch.tensor([0, n]), c["r"], c["cap"], c["loop"],
                         cell=c["cell"][None], pbc=c["pbc"][None])
    _check_single(c, g)


@pytest.mark.gpu
@pytest.mark.parametrize("r,cap", [(2.0, 100000), (1.5, 3), (2.5, 100000)])
def test_product_retry_batched(r, cap, libhgk):
    from hydragnn.ops import radius_graph_pbc

    pos, ptr, cell, pbc, want = _batched(BATCH, False, r, cap)
    np.random.seed(99)
    g = radius_graph_pbc(pos.to(DEV), ptr, r, cap, False, cell=cell, pbc=pbc)
    _check_batched(g, ptr, want)
