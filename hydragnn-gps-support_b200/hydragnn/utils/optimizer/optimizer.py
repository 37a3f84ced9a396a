"""Optimizer selection by name (reference: hydragnn/utils/optimizer/optimizer.py:104-113).
Stock torch optimizers; ZeRO / DeepSpeed variants are outside the B200 hot path."""
import torch

_TABLE = {
    "SGD": torch.optim.SGD, "Adam": torch.optim.Adam, "Adadelta": torch.optim.Adadelta,
    "Adagrad": torch.optim.Adagrad, "Adamax": torch.optim.Adamax, "AdamW": torch.optim.AdamW,
    "RMSprop": torch.optim.RMSprop,
}


def select_optimizer(model, config):
    kind = config["type"]
    if config.get("use_zero_redundancy"):
        raise NotImplementedError("ZeroRedundancyOptimizer is not part of the B200 path (models are a few MB)")
    if kind not in _TABLE:
        raise NameError("The string used to identify the optimizer is NOT recognized by the code")
    params = list(model.parameters())
    kwargs = {}
    if kind in ("Adam", "AdamW", "SGD") and params and all(p.is_cuda for p in params):
        kwargs["fused"] = True      # one kernel for all parameters instead of a dozen foreach launches
    return _TABLE[kind](params, lr=config["learning_rate"], **kwargs)
