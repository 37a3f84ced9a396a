"""Helpers shared by the GPU parity tests and __graft_entry__.smoke().

A "golden case" (tests/golden/reference_fixtures.pt, written by tests/golden/make_golden.py
from the REAL reference modules) holds a seeded mini-batch, the reference model's state dict
and the loss / predictions / parameter gradients / forces the reference produced on CPU.
`run_golden_case` builds the PRODUCT model (hydragnn.models.*Stack on libhgk.so), loads the
same state dict, runs the same step on the GPU and returns both results.
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "hydragnn-gps-support_b200"), os.path.join(ROOT, "tests", "golden")):
    if p not in sys.path:
        sys.path.insert(0, p)

_GOLDEN = None

# fp32 tolerance of BASELINE.json's north_star: rel 1e-5 on losses / energies / forces.
# Parameter gradients pass through one more GEMM chain: 1e-4 as in tests/test_oracle_golden.py.
TOL_VALUE = dict(rtol=1e-5, atol=2e-6)
TOL_GRAD = dict(rtol=1e-4, atol=2e-6)


def golden():
    global _GOLDEN
    if _GOLDEN is None:
        _GOLDEN = torch.load(os.path.join(ROOT, "tests", "golden", "reference_fixtures.pt"),
                             weights_only=False)
    return _GOLDEN


def product_classes():
    from hydragnn.models import EGCLStack, PAINNStack, PNAStack, SCFStack

    return {"SCFStack": SCFStack, "EGCLStack": EGCLStack, "PAINNStack": PAINNStack, "PNAStack": PNAStack}


def to_device(batch, device):
    from hydragnn.graph import Data

    fields = {k: (v.clone().to(device) if torch.is_tensor(v) else v) for k, v in batch.items()}
    return Data(**fields)


def run_step(model, data, cfg, ref_out, device):
    out = {}
    if cfg.get("forces"):
        data.pos.requires_grad_(True)
        pred = model(data)
        loss, _ = model.energy_force_loss(pred, data)[:2]
        out["pred0"] = pred[0].detach().cpu()
    else:
        pred = model(data)
        y = ref_out["y"].to(device)
        head_index = [h.to(device) for h in ref_out["head_index"]]
        loss, _ = model.loss(pred, y, head_index)
        for k, p in enumerate(pred):
            out[f"pred{k}"] = p.detach().cpu()
    loss.backward()
    out["loss"] = loss.detach().cpu()
    out["grads"] = {n: p.grad.detach().cpu() for n, p in model.named_parameters() if p.grad is not None}
    if cfg.get("forces"):
        out["pos_grad"] = data.pos.grad.detach().cpu()
    return out


def run_golden_case(name, device="cuda:0"):
    import make_golden as mg

    fx = golden()[name]
    kind, cfg = fx["kind"], fx["cfg"]
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    model = mg.build(product_classes(), kind, cfg)
    model.load_state_dict(fx["state"], strict=True)
    model = model.to(device)
    if cfg.get("eval"):
        model.eval()
    data = to_device(fx["batch"], device)
    out = run_step(model, data, cfg, fx["out"], device)
    return {"name": name, "cfg": cfg, "out": out, "ref": fx["out"]}


def assert_case_close(res):
    out, ref, cfg = res["out"], res["ref"], res["cfg"]
    torch.testing.assert_close(out["loss"], ref["loss"], **TOL_VALUE)
    k = 0
    while f"pred{k}" in ref:
        torch.testing.assert_close(out[f"pred{k}"], ref[f"pred{k}"], **TOL_VALUE)
        k += 1
    assert set(out["grads"]) == set(ref["grads"]), set(out["grads"]) ^ set(ref["grads"])
    for n, g in ref["grads"].items():
        torch.testing.assert_close(out["grads"][n], g, msg=lambda m, n=n: f"grad {n}: {m}", **TOL_GRAD)
    if cfg.get("forces"):
        torch.testing.assert_close(out["pos_grad"], ref["pos_grad"], **TOL_GRAD)
