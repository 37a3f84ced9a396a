"""hydragnn.run_prediction(config) -> (error, error_rmse_task, true_values, predicted_values), as the
reference (hydragnn/run_prediction.py:34-107): rebuild the loaders and the model from the config,
load ./logs/<log_name>/<log_name>.pk and evaluate the test split."""
import json
import os
from functools import singledispatch

import torch

from .models.create import create_model_config
from .preprocess.load_data import dataset_loading_and_splitting
from .train.train_validate_test import test
from .utils.distributed import get_distributed_model, setup_ddp
from .utils.input_config_parsing.config_utils import get_log_name_config, update_config
from .utils.model import load_existing_model


@singledispatch
def run_prediction(config, use_deepspeed=False):
    raise TypeError("Input must be filename string or configuration dictionary.")


@run_prediction.register
def _(config_file: str, use_deepspeed=False):
    with open(config_file, "r") as fh:
        config = json.load(fh)
    return run_prediction(config, use_deepspeed)


@run_prediction.register
def _(config: dict, use_deepspeed=False):
    if use_deepspeed:
        raise NotImplementedError("deepspeed is not part of the B200 path")
    os.environ.setdefault("SERIALIZED_DATA_PATH", os.getcwd())
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    setup_ddp()
    train_loader, val_loader, test_loader = dataset_loading_and_splitting(config=config)
    config = update_config(config, train_loader, val_loader, test_loader)
    model = create_model_config(config=config["NeuralNetwork"], verbosity=config["Verbosity"]["level"])
    model = get_distributed_model(model, config["Verbosity"]["level"])
    load_existing_model(model, get_log_name_config(config), path="./logs/")
    error, error_rmse_task, true_values, predicted_values = test(
        test_loader, model, config["Verbosity"]["level"],
        compute_grad_energy=config["NeuralNetwork"]["Training"].get("compute_grad_energy", False))
    var = config["NeuralNetwork"]["Variables_of_interest"]
    if var.get("denormalize_output"):
        for k, (lo, hi) in enumerate(var["y_minmax"]):
            true_values[k] = true_values[k] * (hi - lo) + lo
            predicted_values[k] = predicted_values[k] * (hi - lo) + lo
    return error, error_rmse_task, true_values, predicted_values
