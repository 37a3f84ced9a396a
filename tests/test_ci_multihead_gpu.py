"""`-m gpu`: BASELINE config 1 as written -- the reference's tests/inputs/ci_multihead.json (PNA, 2 conv
layers, one graph head + three node heads, task weights 20/1/1/1) on the deterministic toy data:

  * the first three optimisation steps of the product (GPU) against the CPU oracle with the same
    initial weights, batches and optimizer, loss by loss (config 1);
  * hydragnn.run_training + run_prediction end to end with the reference's gates on every head
    (reference tests/test_graphs.py:143-198);
  * hydragnn.run_prediction on a checkpoint written in the REFERENCE's layout by a model with the
    reference's module structure (oracle), reference tests/test_model_loadpred.py:19-98 and
    hydragnn/utils/model/model.py:63-149.
"""
import copy
import os
from collections import OrderedDict

import pytest
import torch

import toy_data

pytestmark = pytest.mark.gpu

CI_MULTIHEAD = {
    "Verbosity": {"level": 0},
    "Dataset": {
        "name": "unit_test_multihead", "format": "unit_test", "compositional_stratified_splitting": True,
        "rotational_invariance": False, "path": {"total": "dataset/unit_test_multihead"},
        "node_features": {"name": ["x", "x2", "x3"], "dim": [1, 1, 1], "column_index": [0, 6, 7]},
        "graph_features": {"name": ["sum_x_x2_x3"], "dim": [1], "column_index": [0]},
    },
    "NeuralNetwork": {
        "Architecture": {
            "global_attn_engine": "", "global_attn_type": "", "mpnn_type": "PNA", "radius": 2.0,
            "max_neighbours": 100, "num_gaussians": 50, "envelope_exponent": 5, "num_radial": 6,
            "num_filters": 126, "periodic_boundary_conditions": False, "pe_dim": 1, "global_attn_heads": 8,
            "hidden_dim": 8, "num_conv_layers": 2,
            "output_heads": {"graph": {"num_sharedlayers": 2, "dim_sharedlayers": 10, "num_headlayers": 2,
                                       "dim_headlayers": [10, 10]},
                             "node": {"num_headlayers": 2, "dim_headlayers": [10, 10], "type": "mlp"}},
            "task_weights": [20.0, 1.0, 1.0, 1.0],
        },
        "Variables_of_interest": {"input_node_features": [0], "output_names": ["sum_x_x2_x3", "x", "x2", "x3"],
                                  "output_index": [0, 0, 1, 2], "type": ["graph", "node", "node", "node"],
                                  "denormalize_output": False},
        "Training": {"num_epoch": 100, "Checkpoint": True, "checkpoint_warmup": 10, "perc_train": 0.7,
                     "loss_function_type": "mse", "batch_size": 16,
                     "Optimizer": {"type": "AdamW", "use_zero_redundancy": False, "learning_rate": 0.01}},
    },
    "Visualization": {"plot_init_solution": False, "plot_hist_solution": False, "create_plots": False},
}


def _setup(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    monkeypatch.setenv("SERIALIZED_DATA_PATH", str(tmp_path))
    toy_data.generate(os.path.join(tmp_path, "dataset", "unit_test_multihead"), 500, 1)
    return copy.deepcopy(CI_MULTIHEAD)


def _loaders_and_config(config):
    from hydragnn.preprocess.load_data import dataset_loading_and_splitting
    from hydragnn.utils.distributed import setup_ddp
    from hydragnn.utils.input_config_parsing.config_utils import update_config

    setup_ddp()
    loaders = dataset_loading_and_splitting(config=config)
    return loaders, update_config(config, *loaders)


def _oracle_model(nn_config):
    """The reference-structured model (oracle) for the same NeuralNetwork config."""
    from hydragnn.models.create import _ARCH_KEYS, _ARCH_KEYS_TAIL
    from oracle import hydragnn_ref as ora

    arch = nn_config["Architecture"]
    kw = {k: arch.get(k) for k in _ARCH_KEYS + _ARCH_KEYS_TAIL}
    kw["freeze_conv"] = kw.pop("freeze_conv_layers")
    return ora.create_model(loss_function_type=nn_config["Training"]["loss_function_type"], **kw)


def _as_oracle_data(batch):
    import make_golden as mg

    return mg.Data(**{k: (v.detach().cpu().clone() if torch.is_tensor(v) else v) for k, v in batch.__dict__.items()
                      if not k.startswith("_")})


def test_first_training_steps_match_oracle(libhgk, tmp_path, monkeypatch):
    from hydragnn.models.create import create_model_config
    from hydragnn.train.train_validate_test import get_head_indices, train_step

    config = _setup(tmp_path, monkeypatch)
    (train_loader, _, _), config = _loaders_and_config(config)
    nn_config = config["NeuralNetwork"]
    product = create_model_config(config=nn_config, verbosity=0)
    oracle = _oracle_model(nn_config)
    oracle.load_state_dict({k: v.detach().cpu() for k, v in product.state_dict().items()}, strict=True)
    product.train(), oracle.train()
    lr = nn_config["Training"]["Optimizer"]["learning_rate"]
    opt_p = torch.optim.AdamW(product.parameters(), lr=lr)
    opt_o = torch.optim.AdamW(oracle.parameters(), lr=lr)
    device = next(product.parameters()).device
    assert device.type == "cuda"
    batches = [b for _, b in zip(range(3), train_loader)]
    assert len(batches) == 3
    for step, batch in enumerate(batches):
        loss_p, tasks_p = train_step(product, opt_p, batch.to(device))
        data = _as_oracle_data(batch)
        head_index = [h.cpu() for h in get_head_indices(product, batch)]
        opt_o.zero_grad()
        loss_o, tasks_o = oracle.loss(oracle(data), data.y, head_index)
        loss_o.backward()
        opt_o.step()
        # step 0: identical weights -> fp32 rel 1e-5.  Later steps start from weights that have each
        # taken an AdamW step (g / sqrt(v) amplifies rounding differences of small gradients): 1e-4.
        tol = 1e-5 if step == 0 else 1e-4
        torch.testing.assert_close(loss_p.cpu(), loss_o.detach(), rtol=tol, atol=1e-6,
                                   msg=lambda m, step=step: f"step {step}: {m}")
        for k, (tp, to) in enumerate(zip(tasks_p, tasks_o)):
            torch.testing.assert_close(tp.detach().cpu(), to.detach(), rtol=tol, atol=1e-6,
                                       msg=lambda m, k=k, step=step: f"step {step} task {k}: {m}")


def test_ci_multihead_training_then_prediction(libhgk, tmp_path, monkeypatch):
    import hydragnn

    config = _setup(tmp_path, monkeypatch)
    hydragnn.run_training(copy.deepcopy(config))
    error, task_errors, true_values, predicted = hydragnn.run_prediction(copy.deepcopy(config))
    assert len(true_values) == 4
    assert error < 0.20, error                               # total RMSE gate of PNA (test_graphs.py:143-158)
    for k in range(4):
        assert true_values[k].shape == predicted[k].shape
        assert float(task_errors[k]) < 0.20, (k, float(task_errors[k]))
        assert torch.mean(torch.abs(true_values[k] - predicted[k])).item() < 0.20, k


def test_prediction_from_reference_layout_checkpoint(libhgk, tmp_path, monkeypatch):
    """A `.pk` as the reference writes it -- {"model_state_dict": {"module." + name: tensor},
    "optimizer_state_dict": ...} from a model with the reference's module tree -- goes through
    hydragnn.run_prediction; the predictions equal that model's own (CPU) predictions."""
    import hydragnn
    from hydragnn.utils.input_config_parsing.config_utils import get_log_name_config

    config = _setup(tmp_path, monkeypatch)
    (_, _, test_loader), full = _loaders_and_config(copy.deepcopy(config))
    oracle = _oracle_model(full["NeuralNetwork"])
    torch.manual_seed(5)
    with torch.no_grad():                                    # not the seeded initial weights, BatchNorm stats moved
        for p in oracle.parameters():
            p.add_(0.05 * torch.randn_like(p))
        for name, buf in oracle.named_buffers():
            if name.endswith("running_mean"):
                buf.add_(0.1 * torch.randn_like(buf))
            elif name.endswith("running_var"):
                buf.mul_(1.0 + 0.2 * torch.rand_like(buf))
    optimizer = torch.optim.AdamW(oracle.parameters(), lr=0.01)
    log_name = get_log_name_config(full)
    os.makedirs(os.path.join("logs", log_name), exist_ok=True)
    torch.save({"model_state_dict": OrderedDict(("module." + k, v) for k, v in oracle.state_dict().items()),
                "optimizer_state_dict": optimizer.state_dict()},
               os.path.join("logs", log_name, log_name + ".pk"))

    error, task_errors, true_values, predicted = hydragnn.run_prediction(copy.deepcopy(config))

    from hydragnn.train.train_validate_test import get_head_indices
    oracle.eval()
    core = type("Core", (), {"num_heads": 4, "head_type": full["NeuralNetwork"]["Variables_of_interest"]["type"],
                             "head_dims": [1, 1, 1, 1]})()
    want = [[] for _ in range(4)]
    with torch.no_grad():
        for batch in test_loader:
            data = _as_oracle_data(batch)
            pred = oracle(data)
            head_index = get_head_indices(core, batch)
            for k in range(4):
                want[k].append(torch.cat([data.y[head_index[k]].reshape(-1, 1), pred[k].reshape(-1, 1)], 1))

    def canon(pairs):
        """The loaders shuffle (also the test split, as in the reference): compare (true, predicted)
        rows as sorted sets."""
        import numpy as np
        a = pairs.double().numpy()
        return torch.from_numpy(a[np.lexsort((a[:, 1], a[:, 0]))])

    for k in range(4):
        got = canon(torch.cat([true_values[k].cpu(), predicted[k].cpu()], 1))
        torch.testing.assert_close(got, canon(torch.cat(want[k])), rtol=1e-5, atol=2e-6)
