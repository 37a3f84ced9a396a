"""Selectors and segment helpers with the reference's names
(hydragnn/utils/model/model.py:29-62, 152-187, 279-286)."""
import os
from collections import OrderedDict

import torch

from ...graph import topo_of
from ...ops import segment_mean, segment_sum


def activation_function_selection(activation_function_string):
    table = {
        "relu": torch.nn.ReLU, "selu": torch.nn.SELU, "prelu": torch.nn.PReLU, "elu": torch.nn.ELU,
        "lrelu_01": lambda: torch.nn.LeakyReLU(0.1), "lrelu_025": lambda: torch.nn.LeakyReLU(0.25),
        "lrelu_05": lambda: torch.nn.LeakyReLU(0.5), "sigmoid": torch.nn.Sigmoid,
    }
    return table[activation_function_string]() if activation_function_string in table else None


def loss_function_selection(loss_function_string):
    if loss_function_string == "mse":
        return torch.nn.functional.mse_loss
    if loss_function_string == "mae":
        return torch.nn.functional.l1_loss
    if loss_function_string == "smooth_l1":
        return torch.nn.SmoothL1Loss
    if loss_function_string == "rmse":
        return lambda x, y: torch.sqrt(torch.nn.functional.mse_loss(x, y))
    if loss_function_string == "GaussianNLLLoss":
        return torch.nn.GaussianNLLLoss()
    raise ValueError(f"Unknown loss function: {loss_function_string}")


def unsorted_segment_mean(data, segment_ids, num_segments, topo=None, side=0):
    """Mean of data rows per segment; `topo` carries the CSR when the caller has one."""
    if topo is None:
        from ...ops.topo import Topo
        topo = Topo(segment_ids.to(torch.int32).contiguous(), None, num_segments)
        side = 0
    return segment_mean(data, topo, side)


def unsorted_segment_sum(data, segment_ids, num_segments, topo=None, side=0):
    if topo is None:
        from ...ops.topo import Topo
        topo = Topo(segment_ids.to(torch.int32).contiguous(), None, num_segments)
        side = 0
    return segment_sum(data, topo, side)


def update_multibranch_heads(output_heads):
    """Wrap legacy single-dict head configs into [{"type": "branch-0", "architecture": ...}]."""
    updated = dict(output_heads)
    for name, val in output_heads.items():
        if isinstance(val, list):
            for branch in val:
                if not (isinstance(branch, dict) and "type" in branch and "architecture" in branch):
                    raise ValueError(f"output_heads['{name}'] does not contain proper branch config, {val}.")
        elif isinstance(val, dict):
            updated[name] = [{"type": "branch-0", "architecture": val}]
        else:
            raise ValueError("Unknown output_heads config!")
    return updated


# ---------------------------------------------------------------- checkpoints ---
def _checkpoint_file(name, path, epoch=None):
    stem = name if epoch is None else f"{name}_epoch_{epoch}"
    return os.path.join(path, name, stem + ".pk")


def save_model(model, optimizer, name, path="./logs/", use_deepspeed=False):
    """Rank 0 writes {"model_state_dict", "optimizer_state_dict"} to ./logs/<name>/<name>.pk
    (the reference's file layout, utils/model/model.py:63-106; HYDRAGNN_EPOCH adds a per-epoch
    file plus a `<name>.pk` symlink).  Keys carry the `module.` prefix of a wrapped model."""
    from ..distributed import get_comm_size_and_rank

    if get_comm_size_and_rank()[1] != 0:
        return
    if hasattr(optimizer, "consolidate_state_dict"):
        optimizer.consolidate_state_dict()
    epoch = os.getenv("HYDRAGNN_EPOCH")
    target = _checkpoint_file(name, path, epoch)
    os.makedirs(os.path.dirname(target), exist_ok=True)
    state = model.state_dict()
    if not next(iter(state), "module.").startswith("module."):
        state = OrderedDict(("module." + k, v) for k, v in state.items())
    torch.save({"model_state_dict": state, "optimizer_state_dict": optimizer.state_dict()}, target)
    if epoch is not None:
        link = _checkpoint_file(name, path)
        if os.path.lexists(link):
            os.remove(link)
        os.symlink(os.path.basename(target), link)


def load_existing_model(model, model_name, path="./logs/", optimizer=None, use_deepspeed=False):
    """Inverse of save_model; accepts checkpoints written with or without the `module.` prefix,
    into a wrapped or a bare model."""
    checkpoint = torch.load(_checkpoint_file(model_name, path), map_location="cpu", weights_only=False)
    state = checkpoint["model_state_dict"]
    wants_prefix = hasattr(model, "module")
    has_prefix = next(iter(state), "").startswith("module.")
    if wants_prefix and not has_prefix:
        state = OrderedDict(("module." + k, v) for k, v in state.items())
    elif has_prefix and not wants_prefix:
        state = OrderedDict((k[len("module."):], v) for k, v in state.items())
    model.load_state_dict(state)
    if optimizer is not None and "optimizer_state_dict" in checkpoint:
        optimizer.load_state_dict(checkpoint["optimizer_state_dict"])


def load_existing_model_config(model, config, path="./logs/", optimizer=None, use_deepspeed=False):
    if config.get("continue"):
        load_existing_model(model, config["startfrom"], path, optimizer)


def get_summary_writer(name, path="./logs/"):
    from ..distributed import get_comm_size_and_rank

    if get_comm_size_and_rank()[1] != 0:
        return None
    try:
        from torch.utils.tensorboard import SummaryWriter
    except Exception:
        return None
    return SummaryWriter(os.path.join(path, name))


class EarlyStopping:
    """Stop when the validation loss has not improved for `patience` epochs."""

    def __init__(self, patience=10, min_delta=0.0):
        self.patience, self.min_delta = patience, min_delta
        self.val_loss_min, self.count = float("inf"), 0

    def __call__(self, val_loss):
        if val_loss > self.val_loss_min + self.min_delta:
            self.count += 1
            return self.count >= self.patience
        self.val_loss_min, self.count = val_loss, 0
        return False


class Checkpoint:
    """Save model + optimizer whenever the metric improves, after `warmup` epochs."""

    def __init__(self, name, warmup=0, path="./logs/", use_deepspeed=False):
        self.name, self.warmup, self.path = name, warmup, path
        self.count, self.min_perf_metric, self.min_delta = 1, float("inf"), 0

    def __call__(self, model, optimizer, perf_metric):
        if perf_metric > self.min_perf_metric + self.min_delta or self.count < self.warmup:
            self.count += 1
            return False
        self.min_perf_metric = perf_metric
        save_model(model, optimizer, name=self.name, path=self.path)
        return True
