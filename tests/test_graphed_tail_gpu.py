"""CUDA-graph replay of the GPS layers' node-level parts (hydragnn/globalAtt/graphed_tail.py) against
the eager path: same weights, same batch, dropout off (the capture's warm-up advances the random
stream, so dropout masks differ between the two by construction), several SGD steps."""
import copy

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _batch(workload, n_graphs, **kwargs):
    from hydragnn.graph import Data
    from hydragnn.ops import radius_graph
    from hydragnn.utils import synthetic as syn

    spec = syn.WORKLOADS[workload]
    rng = np.random.default_rng(5)
    graphs = [spec["sample"](rng, **kwargs) for _ in range(n_graphs)]
    batch = syn.collate(graphs)
    if spec["edges"] == "radius":
        g = radius_graph(batch["pos"].cuda(), batch["ptr"], spec["radius"], spec["max_neighbours"], loop=False,
                         cell=batch.get("cell") if spec["pbc"] else None, pbc=batch.get("pbc") if spec["pbc"] else None)
        batch["edge_index"] = g["edge_index"].cpu()
        if spec["pbc"]:
            batch["edge_shifts"] = g["edge_shifts"].cpu()
    batch = syn.finish_edges(batch)
    dev = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in batch.items() if k not in ("cell", "pbc")}
    return spec, lambda: Data(**{k: (v.clone() if torch.is_tensor(v) else v) for k, v in dev.items()})


def _no_dropout(model):
    from hydragnn.globalAtt.gps import GPSConv

    for m in model.modules():
        if isinstance(m, GPSConv):
            m.dropout = 0.0
        if isinstance(m, torch.nn.Dropout):
            m.p = 0.0


@pytest.mark.parametrize("workload,kwargs", [("schnet_gps_fesi1024", dict(cells=3)), ("egnn_gps_zinc", {})])
def test_graphed_tail_matches_eager(libhgk, workload, kwargs, monkeypatch):
    from hydragnn.globalAtt import graphed_tail
    from hydragnn.models import create_model
    from hydragnn.train import train_step
    from hydragnn.utils.distributed import get_distributed_model

    torch.backends.cuda.matmul.allow_tf32 = False
    spec, fresh = _batch(workload, 3, **kwargs)
    arch = dict(spec["arch"], num_conv_layers=2)
    torch.manual_seed(0)
    base = create_model(**arch).cuda()
    _no_dropout(base)
    results = []
    for enabled in (False, True):
        monkeypatch.setattr(graphed_tail, "ENABLED", enabled)
        monkeypatch.setattr(graphed_tail, "CAPTURE_AFTER", 2)
        model = get_distributed_model(copy.deepcopy(base))
        model.train()
        opt = torch.optim.SGD(model.module.parameters(), lr=1e-2)
        losses = []
        for _ in range(6):
            loss, _ = train_step(model, opt, fresh(), compute_grad_energy=False)
            losses.append(loss.clone())
        graphed = sum(len(m.__dict__.get("_graphed_tails", {})) for m in model.module.modules())
        assert (graphed > 0) == enabled                # the graphs were actually captured / never built
        results.append((torch.stack(losses), {k: v.clone() for k, v in model.module.state_dict().items()}))
    (l0, s0), (l1, s1) = results
    torch.testing.assert_close(l1, l0, rtol=2e-5, atol=1e-7)
    assert set(s0) == set(s1)                          # no extra state_dict keys from the graph holder
    for k in s0:
        if s0[k].dtype.is_floating_point:
            torch.testing.assert_close(s1[k], s0[k], rtol=2e-4, atol=2e-6, msg=lambda m: f"{k}: {m}")
        else:
            assert torch.equal(s0[k], s1[k]), k        # num_batches_tracked: warm-up passes not counted


def test_graphed_tail_steps_aside_for_force_loss(libhgk, monkeypatch):
    """Inside `graphed_tail.eager()` (Base.forward enters it when pos requires grad, i.e. when the
    force loss will differentiate the backward pass) nothing is captured."""
    from hydragnn.globalAtt import graphed_tail
    from hydragnn.globalAtt.gps import GPSConv

    monkeypatch.setattr(graphed_tail, "ENABLED", True)
    monkeypatch.setattr(graphed_tail, "CAPTURE_AFTER", 1)
    layer = GPSConv(128, None, heads=4, dropout=0.0).cuda().train()
    x = torch.randn(256, 128, device="cuda", requires_grad=True)
    with graphed_tail.eager():
        out, _ = layer(x, None)
    assert "_graphed_tails" not in layer.__dict__ or not any(
        e["graphed"] is not None for e in layer.__dict__["_graphed_tails"].values())
    out.sum().backward()
    assert x.grad is not None
