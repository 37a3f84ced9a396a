"""`-m gpu`: the public API end to end on the reference's CI toy problem (BASELINE config 1:
tests/inputs/ci.json -- hidden 8, 2 conv layers, graph head) for every stack on the B200 path:
hydragnn.run_training then hydragnn.run_prediction, with the reference's own acceptance gates
(reference tests/test_graphs.py:143-198)."""
import copy
import os

import pytest
import torch

import toy_data

pytestmark = pytest.mark.gpu

CI_CONFIG = {
    "Verbosity": {"level": 0},
    "Dataset": {
        "name": "unit_test_singlehead", "format": "unit_test", "compositional_stratified_splitting": True,
        "rotational_invariance": False,
        "path": {"train": "dataset/unit_test_singlehead_train", "test": "dataset/unit_test_singlehead_test",
                 "validate": "dataset/unit_test_singlehead_validate"},
        "node_features": {"name": ["x", "x2", "x3"], "dim": [1, 1, 1], "column_index": [0, 6, 7]},
        "graph_features": {"name": ["sum_x_x2_x3"], "dim": [1], "column_index": [0]},
    },
    "NeuralNetwork": {
        "Architecture": {
            "global_attn_engine": "", "global_attn_type": "", "mpnn_type": "PNA", "radius": 2.0,
            "max_neighbours": 100, "num_gaussians": 50, "num_filters": 126, "num_radial": 6,
            "envelope_exponent": 5,
            "periodic_boundary_conditions": False, "pe_dim": 1, "global_attn_heads": 8, "hidden_dim": 8,
            "num_conv_layers": 2,
            "output_heads": {"graph": {"num_sharedlayers": 2, "dim_sharedlayers": 4, "num_headlayers": 2,
                                       "dim_headlayers": [10, 10]},
                             "node": {"num_headlayers": 2, "dim_headlayers": [4, 4], "type": "mlp"}},
            "task_weights": [1.0],
        },
        "Variables_of_interest": {"input_node_features": [0], "output_names": ["sum_x_x2_x3"],
                                  "output_index": [0], "type": ["graph"], "denormalize_output": False},
        "Training": {"num_epoch": 60, "perc_train": 0.7, "EarlyStopping": True, "patience": 10,
                     "Checkpoint": True, "checkpoint_warmup": 10, "loss_function_type": "mse", "batch_size": 32,
                     "Optimizer": {"type": "AdamW", "use_zero_redundancy": False, "learning_rate": 0.02}},
    },
    "Visualization": {"plot_init_solution": False, "plot_hist_solution": False, "create_plots": False},
}


# RMSE / MAE gates of the reference (tests/test_graphs.py:143-158)
GATES = {"PNA": (0.20, 0.20), "SchNet": (0.20, 0.20), "EGNN": (0.20, 0.20), "PNAPlus": (0.20, 0.20),
         "PNAEq": (0.60, 0.60), "PAINN": (0.60, 0.60)}


@pytest.mark.parametrize("mpnn_type", sorted(GATES))
def test_run_training_then_prediction_on_ci_toy(libhgk, tmp_path, monkeypatch, mpnn_type):
    import hydragnn

    monkeypatch.chdir(tmp_path)
    monkeypatch.setenv("SERIALIZED_DATA_PATH", str(tmp_path))
    for split, count, seed in (("train", 350, 1), ("test", 75, 2), ("validate", 75, 3)):
        toy_data.generate(os.path.join(tmp_path, "dataset", f"unit_test_singlehead_{split}"), count, seed)
    config = copy.deepcopy(CI_CONFIG)
    config["NeuralNetwork"]["Architecture"]["mpnn_type"] = mpnn_type
    hydragnn.run_training(copy.deepcopy(config))
    log_dirs = os.listdir(tmp_path / "logs")
    assert len(log_dirs) == 1 and os.path.exists(tmp_path / "logs" / log_dirs[0] / (log_dirs[0] + ".pk"))
    error, task_errors, true_values, predicted = hydragnn.run_prediction(copy.deepcopy(config))
    assert true_values[0].shape == predicted[0].shape and true_values[0].shape[0] == 75
    rmse = torch.sqrt(torch.mean((true_values[0] - predicted[0]) ** 2)).item()
    mae = torch.mean(torch.abs(true_values[0] - predicted[0])).item()
    assert rmse < GATES[mpnn_type][0] and mae < GATES[mpnn_type][1], (mpnn_type, rmse, mae)
