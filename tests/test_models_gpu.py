"""`-m gpu`: the product stacks (hydragnn.models.*Stack on libhgk.so) on the B200 against the golden
outputs of the REAL reference modules (tests/golden/reference_fixtures.pt): same state dict, same
seeded batch; loss / predictions within fp32 rel 1e-5, parameter gradients and forces within 1e-4."""
import pytest
import torch

import parity_util as pu

pytestmark = pytest.mark.gpu

CASES = ["schnet_plain", "schnet_equiv", "schnet_gps", "egnn_gps_edge", "egnn_equiv", "painn_forces",
         "schnet_forces", "egnn_forces", "pna_multihead", "pna_edge"]


@pytest.mark.parametrize("name", CASES)
def test_product_stack_matches_reference(libhgk, name):
    res = pu.run_golden_case(name, device="cuda:0")
    pu.assert_case_close(res)


@pytest.mark.parametrize("name", ["schnet_plain", "painn_forces", "egnn_gps_edge"])
def test_step_is_bit_reproducible(libhgk, name):
    """Deterministic segmented reductions: two runs of the same step give identical bits."""
    a = pu.run_golden_case(name, device="cuda:0")["out"]
    b = pu.run_golden_case(name, device="cuda:0")["out"]
    assert torch.equal(a["loss"], b["loss"])
    for n in a["grads"]:
        assert torch.equal(a["grads"][n], b["grads"][n]), n
