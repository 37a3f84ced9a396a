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
