"""Selectors and segment helpers with the reference's names
(hydragnn/utils/model/model.py:29-62, 152-187, 279-286)."""
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
