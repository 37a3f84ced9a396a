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
            g = radius_graph(batch["pos"].cuda(), batch["ptr"], spec["radius"], spec["max_neighbours"], loop=False,
                             cell=batch.get("cell") if spec["pbc"] else None, pbc=batch.get("pbc") if spec["pbc"] else None)
            batch["edge_index"] = g["edge_index"].cpu()
            if spec["pbc"]:
                batch["edge_shifts"] = g["edge_shifts"].cpu()
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


def _setup(workload):
    from hydragnn.models import create_model

    torch.backends.cuda.matmul.allow_tf32 = False
    # eight (force workload: three) graphs per batch, every batch a different draw: N and E vary inside and
    # across buckets
    spec, batches = _batches(workload, [3] * 6 if workload == "painn_oc20_forces" else [8] * 6)
    arch = dict(spec["arch"], num_conv_layers=2)
    torch.manual_seed(0)
    base = create_model(**arch).cuda()
    _no_dropout(base)
    base.train()
    return base, batches


def _runner(base, use_graph, lr, forces=False):
    from hydragnn.train import train_step
    from hydragnn.train.step_graph import BucketedStepGraph
    from hydragnn.utils.distributed import get_distributed_model

    model = get_distributed_model(copy.deepcopy(base))
    opt = torch.optim.AdamW(model.module.parameters(), lr=lr, fused=True, capturable=True)
    sg = BucketedStepGraph(model, opt, node_bucket=32, edge_bucket=256) if use_graph else None

    def step(data):
        if use_graph:
            assert sg.eligible(data, forces), "step graph not eligible"
            return sg.step(data, compute_grad_energy=forces)[0]
        return train_step(model, opt, data, compute_grad_energy=forces)[0]
    return model, sg, step


@pytest.mark.parametrize("workload", ["schnet_qm9", "egnn_gps_zinc", "painn_oc20_forces"])
def test_step_graph_gradients_match_eager(libhgk, workload):
    """Learning rate 0 (the weights stay put): every batch, first through a freshly captured graph and again
    through its replay, must give the eager step's loss (rel 1e-5) and gradients.  The PaiNN case runs the
    energy + force loss: the captured step contains the double backward."""
    forces = workload == "painn_oc20_forces"
    base, batches = _setup(workload)
    model_e, _, step_e = _runner(base, False, 0.0, forces)
    model_g, sg, step_g = _runner(base, True, 0.0, forces)
    for rep in range(2):
        for make in batches:
            le, lg = float(step_e(make())), float(step_g(make()))
            assert abs(lg - le) <= 1e-5 * abs(le) + 1e-7, (rep, le, lg)
            ge = {n: p.grad.detach().clone() for n, p in model_e.module.named_parameters() if p.grad is not None}
            gg = {n: p.grad.detach().clone() for n, p in model_g.module.named_parameters() if p.grad is not None}
            assert set(ge) == set(gg)
            floor = 1e-6 * max(float(g.abs().max()) for g in ge.values())
            for n, g in ge.items():
                scale = float(g.abs().max())
                torch.testing.assert_close(gg[n], g, rtol=1e-4, atol=max(1e-4 * scale, floor), msg=lambda m, n=n: f"{n}: {m}")
    assert sg.replays == 12 and sg.eager_steps == 0
    assert 1 <= sg.captures < len(batches) + 1                      # buckets are shared between batches


@pytest.mark.parametrize("workload", ["schnet_qm9", "egnn_gps_zinc"])
def test_step_graph_training_tracks_eager(libhgk, workload):
    """Real optimisation steps: the two trajectories start together (first losses rel 1e-5) and stay close;
    they may drift apart slowly -- AdamW turns rounding-level differences of near-zero gradients into
    full-size updates -- so the later steps are held to 1 % only."""
    base, batches = _setup(workload)
    _, _, step_e = _runner(base, False, 1e-3)
    _, sg, step_g = _runner(base, True, 1e-3)
    eager, graph = [], []
    for rep in range(2):
        for make in batches:
            eager.append(float(step_e(make())))
            graph.append(float(step_g(make())))
    assert abs(graph[0] - eager[0]) <= 1e-5 * abs(eager[0]) + 1e-7
    np.testing.assert_allclose(graph[:3], eager[:3], rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(graph, eager, rtol=1e-2, atol=1e-5)


def test_run_training_with_step_graphs(libhgk, tmp_path, monkeypatch):
    """hydragnn.run_training / run_prediction on the CI toy with HYDRAGNN_STEP_GRAPH=1 (tiny graphs of 2-8 atoms,
    batch 32: many size buckets, a ragged last batch) reaches the reference's accuracy gate."""
    import os

    import hydragnn
    import toy_data
    from test_run_training_gpu import CI_CONFIG

    monkeypatch.chdir(tmp_path)
    monkeypatch.setenv("SERIALIZED_DATA_PATH", str(tmp_path))
    monkeypatch.setenv("HYDRAGNN_STEP_GRAPH", "1")
    monkeypatch.setenv("HYDRAGNN_STEP_GRAPH_NODES", "32")
    monkeypatch.setenv("HYDRAGNN_STEP_GRAPH_EDGES", "128")
    for split, count, seed in (("train", 350, 1), ("test", 75, 2), ("validate", 75, 3)):
        toy_data.generate(os.path.join(tmp_path, "dataset", f"unit_test_singlehead_{split}"), count, seed)
    config = copy.deepcopy(CI_CONFIG)
    config["NeuralNetwork"]["Architecture"]["mpnn_type"] = "SchNet"
    hydragnn.run_training(copy.deepcopy(config))
    error, task_errors, true_values, predicted = hydragnn.run_prediction(copy.deepcopy(config))
    assert torch.sqrt(torch.mean((true_values[0] - predicted[0]) ** 2)).item() < 0.20


def _rank_worker(rank, world, port, out_dir):
    """One data-parallel rank (both ranks share cuda:0; gloo carries the CUDA all-reduce)."""
    import os
    import sys

    here = os.path.dirname(os.path.abspath(__file__))
    sys.path[:0] = [os.path.join(os.path.dirname(here), "hydragnn-gps-support_b200"), here]
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK="0", HYDRAGNN_BACKEND="gloo")
    import torch.distributed as dist
    from hydragnn.train import train_step
    from hydragnn.utils.distributed import setup_ddp

    setup_ddp()
    base, batches = _setup("schnet_qm9")
    mine = batches[rank::world]                                     # each rank its own shard of the stream
    twin = copy.deepcopy(base)                                      # this rank's own gradients, no exchange
    opt_twin = torch.optim.AdamW(twin.parameters(), lr=0.0, fused=True)
    model, sg, step = _runner(base, True, 0.0)
    assert model.averager.enabled
    record = []
    for rep in range(2):
        for make in mine:
            train_step(twin, opt_twin, make())
            loss = step(make())
            record.append({"loss": float(loss),
                           "local": {n: p.grad.detach().cpu().clone() for n, p in twin.named_parameters()
                                     if p.grad is not None},
                           "avg": {n: p.grad.detach().cpu().clone() for n, p in model.module.named_parameters()
                                   if p.grad is not None}})
    assert sg.eager_steps == 0 and sg.replays == len(record)
    # real steps: the ranks must stay in lockstep
    for g in sg.optimizer.param_groups:
        g["lr"] = 1e-3
    for make in mine:
        step(make())
    params = {n: p.detach().cpu().clone() for n, p in model.module.named_parameters()}
    torch.save({"record": record, "params": params}, os.path.join(out_dir, f"rank{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_step_graph_two_ranks_average_gradients(libhgk, tmp_path):
    """World size 2: every rank replays its own bucket's graph (the ranks' batches differ in size, so do their
    buckets and capture steps), the flat gradient buffer is averaged between the replay and the captured
    optimizer step.  Averaged gradient = mean of the two ranks' own eager gradients; parameters after real
    steps are identical on both ranks."""
    import socket

    import torch.multiprocessing as mp

    def run_ranks():
        with socket.socket() as s:
            s.bind(("127.0.0.1", 0))
            port = s.getsockname()[1]
        mp.spawn(_rank_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)

    try:
        run_ranks()
    except Exception as exc:          # rendezvous / process start-up trouble (two processes share one GPU here): once more
        print(f"two-rank run failed once, retrying: {exc!r}")
        run_ranks()
    r0, r1 = (torch.load(tmp_path / f"rank{r}.pt") for r in range(2))
    assert len(r0["record"]) == len(r1["record"]) == 6
    for a, b in zip(r0["record"], r1["record"]):
        assert set(a["avg"]) == set(a["local"]) == set(b["local"])
        for n in a["avg"]:
            want = 0.5 * (a["local"][n] + b["local"][n])
            scale = float(want.abs().max())
            torch.testing.assert_close(a["avg"][n], b["avg"][n], rtol=0, atol=0, msg=lambda m, n=n: f"{n}: {m}")
            torch.testing.assert_close(a["avg"][n], want, rtol=1e-4, atol=max(1e-4 * scale, 1e-9),
                                       msg=lambda m, n=n: f"{n}: {m}")
    for n in r0["params"]:
        torch.testing.assert_close(r0["params"][n], r1["params"][n], rtol=0, atol=0)
