"""hydragnn.run_training(config) -- same contract as the reference (hydragnn/run_training.py:48-182):
`config` is a JSON file name or a dict; side effects ./logs/<log_name>/{run.log, config.json,
<log_name>.pk}; SERIALIZED_DATA_PATH defaults to the working directory."""
import json
import os
from functools import singledispatch

import torch
import torch.distributed as dist
from torch.optim.lr_scheduler import ReduceLROnPlateau

from .models.create import create_model_config
from .preprocess.load_data import dataset_loading_and_splitting
from .train.train_validate_test import train_validate_test
from .utils.distributed import get_distributed_model, setup_ddp
from .utils.input_config_parsing.config_utils import get_log_name_config, save_config, update_config
from .utils.model import get_summary_writer, load_existing_model_config, save_model
from .utils.optimizer import select_optimizer
from .utils.print import print_distributed, setup_log


@singledispatch
def run_training(config, use_deepspeed=False):
    raise TypeError("Input must be filename string or configuration dictionary.")


@run_training.register
def _(config_file: str, use_deepspeed=False):
    with open(config_file, "r") as fh:
        config = json.load(fh)
    run_training(config, use_deepspeed)


@run_training.register
def _(config: dict, use_deepspeed=False):
    if use_deepspeed:
        raise NotImplementedError("deepspeed is not part of the B200 path")
    os.environ.setdefault("SERIALIZED_DATA_PATH", os.getcwd())
    # fp32 parity with the reference: no TF32 in the surrounding torch GEMMs
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False

    setup_log(get_log_name_config(config))
    setup_ddp()
    train_loader, val_loader, test_loader = dataset_loading_and_splitting(config=config)
    config = update_config(config, train_loader, val_loader, test_loader)
    verbosity = config["Verbosity"]["level"]

    model = create_model_config(config=config["NeuralNetwork"], verbosity=verbosity)
    model = get_distributed_model(model, verbosity,
                                  sync_batch_norm=config["NeuralNetwork"]["Architecture"]["SyncBatchNorm"])
    optimizer = select_optimizer(model, config["NeuralNetwork"]["Training"]["Optimizer"])
    scheduler = ReduceLROnPlateau(optimizer, mode="min", factor=0.5, patience=5, min_lr=0.00001)
    log_name = get_log_name_config(config)
    writer = get_summary_writer(log_name)
    if dist.is_initialized():
        dist.barrier()
    save_config(config, log_name)
    load_existing_model_config(model, config["NeuralNetwork"]["Training"], optimizer=optimizer)
    print_distributed(verbosity, "Starting training with the configuration: \n"
                      + json.dumps(config, indent=4, sort_keys=True))
    vis = config.get("Visualization", {})
    train_validate_test(
        model, optimizer, train_loader, val_loader, test_loader, writer, scheduler,
        config["NeuralNetwork"], log_name, verbosity,
        vis.get("plot_init_solution", False), vis.get("plot_hist_solution", False), vis.get("create_plots", False),
        compute_grad_energy=config["NeuralNetwork"]["Training"]["compute_grad_energy"])
    save_model(model, optimizer, log_name)
