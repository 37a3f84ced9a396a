"""`-m gpu`: whole training steps replayed as CUDA graphs over size buckets (hydragnn/train/step_graph.py)
against the eager `train_step`: same initial weights, the same sequence of mini-batches of DIFFERENT
sizes (several batches per bucket, more than one bucket), dropout off; per-step losses and the
parameters after the last step must agree -- the padding (pad nodes / pad edges / pad graph, masked loss,
BatchNorm over the valid rows only) must be invisible."""
import copy

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _batches(workload, sizes, seed=3):
    from hydragnn.graph import Data
    from hydragnn.utils import synthetic as syn

    spec = syn.WORKLOADS[workload]
    rng = np.random.default_rng(seed)
    out = []
    for n_graphs in sizes:
        graphs = [spec["sample"](rng) for _ in range(n_graphs)]
        batch = syn.collate(graphs)
        if spec["edges"] == "radius":
            from hydragnn.ops import radius_graph
            g = radius_graph(batch["pos"].cuda(), batch["ptr"], spec["radius"], spec["max_neighbours"], loop=False)
            batch["edge_index"] = g["edge_index"].cpu()
        batch = syn.finish_edges(batch)
        dev = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in batch.items() if k not in ("cell", "pbc")}
        out.append(lambda dev=dev: Data(**{k: (v.clone() if torch.is_tensor(v) else v) for k, v in dev.items()}))
    return spec, out


def _no_dropout(model):
    for m in model.modules():
        if isinstance(m, torch.nn.Dropout):
            m.p = 0.0
        elif hasattr(m, "attn") and isinstance(getattr(m, "dropout", None), float):
            m.dropout = 0.0


@pytest.mark.parametrize("workload", ["schnet_qm9", "egnn_gps_zinc"])
def test_step_graph_matches_eager_training(libhgk, workload):
    from hydragnn.models import create_model
    from hydragnn.train import train_step
    from hydragnn.train.step_graph import BucketedStepGraph
    from hydragnn.utils.distributed import get_distributed_model

    torch.backends.cuda.matmul.allow_tf32 = False
    # eight graphs per batch, every batch a different draw: N and E vary inside and across buckets
    spec, batches = _batches(workload, [8, 8, 8, 8, 8, 8])
    arch = dict(spec["arch"], num_conv_layers=2)
    torch.manual_seed(0)
    base = create_model(**arch).cuda()
    _no_dropout(base)
    base.train()

    def run(use_graph):
        model = get_distributed_model(copy.deepcopy(base))
        opt = torch.optim.AdamW(model.module.parameters(), lr=1e-3, fused=True, capturable=True)
        sg = BucketedStepGraph(model, opt, node_bucket=32, edge_bucket=256) if use_graph else None
        losses = []
        for rep in range(2):
            for make in batches:
                data = make()
                if use_graph:
                    assert sg.eligible(data), "step graph not eligible"
                    loss, _ = sg.step(data)
                else:
                    loss, _ = train_step(model, opt, data)
                losses.append(float(loss))
        return losses, {k: v.detach().cpu() for k, v in model.module.state_dict().items()}, sg

    eager_losses, eager_state, _ = run(False)
    graph_losses, graph_state, sg = run(True)
    assert sg.replays == len(graph_losses) and sg.eager_steps == 0
    assert 1 <= sg.captures < len(batches) + 1                      # buckets are shared between batches
    # the first step starts from identical weights: fp32 rel 1e-5; later steps compound rounding
    # differences through AdamW (g / sqrt(v) amplifies them for small gradients)
    assert abs(graph_losses[0] - eager_losses[0]) <= 1e-5 * abs(eager_losses[0]) + 1e-7
    np.testing.assert_allclose(graph_losses, eager_losses, rtol=2e-4, atol=1e-6)
    for k, v in eager_state.items():
        if v.is_floating_point():
            scale = float(v.abs().max())
            torch.testing.assert_close(graph_state[k], v, rtol=1e-3, atol=2e-4 * max(scale, 1e-3), msg=lambda m, k=k: f"{k}: {m}")
        else:
            assert torch.equal(graph_state[k], v), k
