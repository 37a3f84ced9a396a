"""`-m gpu`: the product stacks (hydragnn.models.*Stack on libhgk.so) on the B200 against the golden
outputs of the REAL reference modules (tests/golden/reference_fixtures.pt): same state dict, same
seeded batch; loss / predictions / BatchNorm running statistics within fp32 rel 1e-5; parameter
gradients and forces within 1e-5 of the fp64 oracle relative to the tensor's scale, or as close to it
as the reference's own fp32 result is (parity_util.assert_close_to_truth)."""
import pytest
import torch

import parity_util as pu

pytestmark = pytest.mark.gpu

CASES = ["schnet_plain", "schnet_equiv", "schnet_gps", "egnn_gps_edge", "egnn_equiv", "painn_forces",
         "schnet_forces", "egnn_forces", "pna_multihead", "pna_edge",
         # GPS in train mode (BatchNorm batch statistics), dropout = 0: tests/golden/make_golden_train.py
         "schnet_gps_train", "egnn_gps_edge_train", "painn_gps_train", "schnet_gps_train_wide",
         "egnn_gps_train_wide"]


@pytest.mark.parametrize("name", CASES)
def test_product_stack_matches_reference(libhgk, name):
    res = pu.run_golden_case(name, device="cuda:0")
    # gradients / forces: bounded against the oracle evaluated in fp64 (parity_util, "fp64 truth")
    pu.assert_case_close(res, truth=pu.oracle_truth_case(name))


@pytest.mark.parametrize("name", ["schnet_plain", "painn_forces", "egnn_gps_edge"])
def test_step_is_bit_reproducible(libhgk, name):
    """Deterministic segmented reductions: two runs of the same step give identical bits."""
    a = pu.run_golden_case(name, device="cuda:0")["out"]
    b = pu.run_golden_case(name, device="cuda:0")["out"]
    assert torch.equal(a["loss"], b["loss"])
    for n in a["grads"]:
        assert torch.equal(a["grads"][n], b["grads"][n]), n


def _workload_parity(workload, num_graphs, num_layers, train, sample_kwargs, with_truth=True):
    """Product on the GPU vs the CPU oracle (fp32, and fp64 as the truth) on a synthetic batch of a
    BASELINE workload with identical weights.  train=True: train mode with dropout = 0 (BatchNorm
    batch statistics, the mode bench.py times); otherwise eval."""
    import numpy as np
    from hydragnn.graph import Data
    from hydragnn.models import create_model
    from hydragnn.ops import radius_graph
    from hydragnn.utils import synthetic as syn
    from oracle import hydragnn_ref as ora
    import make_golden as mg

    torch.backends.cuda.matmul.allow_tf32 = False
    spec = syn.WORKLOADS[workload]
    rng = np.random.default_rng(11)
    graphs = [spec["sample"](rng, **sample_kwargs) for _ in range(num_graphs)]
    batch = syn.collate(graphs)
    if spec["edges"] == "radius":
        g = radius_graph(batch["pos"].cuda(), batch["ptr"], spec["radius"], spec["max_neighbours"], loop=False,
                         cell=batch.get("cell") if spec["pbc"] else None, pbc=batch.get("pbc") if spec["pbc"] else None)
        batch["edge_index"] = g["edge_index"].cpu()
        if spec["pbc"]:
            batch["edge_shifts"] = g["edge_shifts"].cpu()
    batch = syn.finish_edges(batch)
    arch = dict(spec["arch"], num_conv_layers=num_layers)

    def oracle_run(dtype):
        with pu.default_dtype(dtype):
            ref = ora.create_model(**arch)
            ref.load_state_dict(state, strict=True) if state else None
            pu.set_dropout(ref, 0.0)
            ref.train() if train else ref.eval()
            data = mg.Data(**{k: ((v.to(dtype) if v.is_floating_point() else v.clone()) if torch.is_tensor(v) else v)
                              for k, v in batch.items()})
            pos = data.pos = data.pos.detach().clone()       # a leaf of this run
            if spec["forces"]:
                pos.requires_grad_(True)
                loss = ref.energy_force_loss(ref(data), data)[0]
            else:
                loss = ref.loss(ref(data), data.y, [torch.arange(data.y.shape[0])])[0]
            loss.backward()
            out = {"loss": loss.detach(), "grads": {n: p.grad for n, p in ref.named_parameters()}}
            if spec["forces"]:
                out["forces"] = -pos.grad.detach()           # d loss / d pos (through the force term too)
        return out, ref

    state = None
    with pu.default_dtype(torch.float32):
        seed_model = ora.create_model(**arch)
    state = {k: v.clone() for k, v in seed_model.state_dict().items()}
    ref32, _ = oracle_run(torch.float32)
    truth = oracle_run(torch.float64)[0] if with_truth else None

    model = create_model(**arch)
    model.load_state_dict(state, strict=True)
    pu.set_dropout(model, 0.0)
    model.train() if train else model.eval()
    dev = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in batch.items() if k not in ("cell", "pbc")}
    d = Data(**dev)
    pos_gpu = d.pos
    if spec["forces"]:
        pos_gpu.requires_grad_(True)
        loss = model.energy_force_loss(model(d), d)[0]
    else:
        loss = model.loss(model(d), d.y, [torch.arange(d.y.shape[0], device="cuda")])[0]
    loss.backward()
    torch.testing.assert_close(loss.detach().cpu(), ref32["loss"], rtol=1e-5, atol=1e-6)
    if spec["forces"]:
        if truth is not None:
            pu.assert_close_to_truth(-pos_gpu.grad, ref32["forces"], truth["forces"], "d loss / d pos")
        else:
            torch.testing.assert_close(-pos_gpu.grad.cpu(), ref32["forces"], rtol=1e-4, atol=1e-6)
    floor = pu.grad_floor(truth["grads"].values()) if truth is not None else 0.0
    for n, p in model.named_parameters():
        g32 = ref32["grads"][n]
        if g32 is None:
            assert p.grad is None or float(p.grad.abs().max()) == 0.0, n
            continue
        if truth is not None:
            pu.assert_close_to_truth(p.grad, g32, truth["grads"][n], f"grad {n}", floor=floor)
        else:
            scale = max(float(g32.abs().max()), 1e-6)
            torch.testing.assert_close(p.grad.cpu(), g32, rtol=1e-4, atol=1e-4 * scale,
                                       msg=lambda m, n=n: f"{n}: {m}")
    return float(loss)


@pytest.mark.parametrize("workload,kwargs,train", [
    ("schnet_gps_fesi1024", dict(cells=3), False), ("schnet_qm9", {}, False), ("egnn_gps_zinc", {}, False),
    ("painn_oc20_forces", {}, False),
    ("schnet_gps_fesi1024", dict(cells=3), True), ("egnn_gps_zinc", {}, True)])
def test_full_width_models_match_cpu_oracle(libhgk, workload, kwargs, train):
    """hidden 128 / 50 gaussians (the BASELINE widths, where the tensor-core filter / attention /
    node-row kernels are selected) on a small batch, eval and train mode: loss rel 1e-5 against the
    fp32 oracle; parameter gradients and forces bounded against the fp64 oracle."""
    _workload_parity(workload, 3, 2, train, kwargs)


def test_config5_full_size_train_step_matches_oracle(libhgk):
    """BASELINE config 5 exactly as bench.py times it: 8 crystals of 1024 atoms (N = 8192, one
    whole-batch attention segment, E ~ 474k), 4 SchNet+GPS layers, train mode (dropout = 0 so that the
    step is deterministic).  One oracle step costs a few seconds of host time.  The fp64 truth run
    needs ~40 GB of host memory for the dense [4, 8192, 8192] attention blocks: taken when the box
    has it, otherwise gradients are compared with the fp32 oracle at 1e-4 of their scale."""
    import psutil

    roomy = psutil.virtual_memory().available > 96 * 2 ** 30
    _workload_parity("schnet_gps_fesi1024", 8, 4, True, {}, with_truth=roomy)
