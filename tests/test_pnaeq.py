"""PNAEq (reference hydragnn/models/PNAEqStack.py; the model of the physics-informed multibranch configuration, with its force loss) against golden outputs of the
REAL reference module (tests/golden/make_golden_pnaeq.py -> pnaeq_fixtures.pt):
  * CPU: the oracle restatement (bit exact), and the product stack's host logic with the C-ABI
    launchers swapped for the oracle kernels;
  * GPU: the product stack on libhgk.so -- loss / predictions rel 1e-5, gradients rel 1e-4."""
import os
import sys

import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_golden as mg  # noqa: E402
import make_golden_pnaeq as mp  # noqa: E402
import parity_util as pu  # noqa: E402
from oracle import hydragnn_ref as ora  # noqa: E402

FX = torch.load(os.path.join(HERE, "golden", "pnaeq_fixtures.pt"), weights_only=False)
CASES = sorted(FX)


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference(name):
    fx = FX[name]
    model = mp.build(ora.PNAEqStack, fx["cfg"])
    model.load_state_dict(fx["state"], strict=True)
    if fx["cfg"].get("eval"):
        model.eval()
    out = mg.run_model(model, fx["batch"], fx["cfg"])
    assert torch.equal(out["loss"], fx["out"]["loss"])
    for n, g in fx["out"]["grads"].items():
        torch.testing.assert_close(out["grads"][n], g, rtol=1e-6, atol=1e-7)


def _run_product(name, device):
    from hydragnn.models import PNAEqStack

    fx = FX[name]
    cfg = fx["cfg"]
    torch.backends.cuda.matmul.allow_tf32 = False
    model = mp.build(PNAEqStack, cfg)
    model.load_state_dict(fx["state"], strict=True)       # same parameter names as the reference
    model = model.to(device)
    if cfg.get("eval"):
        model.eval()
    out = pu.run_step(model, pu.to_device(fx["batch"], device), cfg, fx["out"], device)
    return {"name": name, "cfg": cfg, "out": out, "ref": fx["out"]}


@pytest.mark.parametrize("name", CASES)
def test_product_host_logic_on_cpu(name):
    from cpu_patch import cpu_kernels

    with cpu_kernels():
        res = _run_product(name, "cpu")
    pu.assert_case_close(res)


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_product_matches_reference_on_gpu(libhgk, name):
    pu.assert_case_close(_run_product(name, "cuda:0"))


@pytest.mark.gpu
def test_create_model_accepts_pnaeq(libhgk):
    from hydragnn.models import create_model

    cfg = FX["pnaeq_plain"]["cfg"]
    model = create_model("PNAEq", cfg["input_dim"], cfg["hidden_dim"], cfg["output_dim"], 0, None, None, 0,
                         cfg["output_type"], cfg["heads"], "relu", "mse", cfg["task_weights"], cfg["num_conv_layers"],
                         pna_deg=cfg["pna_deg"], num_radial=cfg["num_radial"], radius=cfg["radius"])
    assert set(model.state_dict()) == set(FX["pnaeq_plain"]["state"])
