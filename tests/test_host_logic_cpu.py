"""Host-side logic of the product stacks (weight splitting, CSR sides, the autograd closure used
by the force loss) against the golden outputs of the REAL reference modules -- on CPU, with the
C-ABI launchers swapped for oracle/kernel_spec.py (tests/cpu_patch.py).  The kernels themselves
are checked by the `-m gpu` tests."""
import pytest
import torch

import parity_util as pu
from cpu_patch import cpu_kernels

CASES = ["schnet_plain", "schnet_equiv", "schnet_gps", "egnn_gps_edge", "egnn_equiv", "painn_forces",
         "schnet_forces", "egnn_forces", "pna_multihead", "pna_edge"]


@pytest.mark.parametrize("name", CASES)
def test_product_stack_matches_reference_on_cpu(name):
    with cpu_kernels():
        res = pu.run_golden_case(name, device="cpu")
    pu.assert_case_close(res)
