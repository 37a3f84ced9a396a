"""JSON config handling: same keys, defaults and derived fields as the reference's
hydragnn/utils/input_config_parsing/config_utils.py (update_config :25-161, edge_dim :175-198,
output dims :219-260, log name :314-349), written as tables instead of if-chains."""
import json
import os
import pickle
from copy import deepcopy

import torch

from ..distributed import get_comm_size_and_rank
from ..model import update_multibranch_heads

# Architecture keys that default to None when absent (create_model reads all of them)
_ARCH_NONE_DEFAULTS = (
    "radius", "radial_type", "distance_transform", "num_gaussians", "num_filters", "envelope_exponent",
    "num_after_skip", "num_before_skip", "basis_emb_size", "int_emb_size", "out_emb_size", "num_radial",
    "num_spherical", "correlation", "max_ell", "node_max_ell",
)
_ARCH_DEFAULTS = {"freeze_conv_layers": False, "initial_bias": None, "activation_function": "relu",
                  "SyncBatchNorm": False}
_EDGE_MODELS = ("GAT", "PNA", "PNAPlus", "PAINN", "PNAEq", "CGCNN", "SchNet", "EGNN", "DimeNet", "MACE")
_EQUIVARIANT_MODELS = ("EGNN", "SchNet", "PNAEq", "PAINN", "MACE")


def _dataset_of(loader):
    ds = loader.dataset
    return getattr(ds, "samples", ds)          # PackedDataset keeps the host samples next to its device arenas


def _graph_size_variable(*loaders):
    env = os.getenv("HYDRAGNN_USE_VARIABLE_GRAPH_SIZE")
    if env is not None:
        return bool(int(env))
    sizes = set()
    for loader in loaders:
        for data in _dataset_of(loader):
            sizes.add(int(data.num_nodes))
            if len(sizes) > 1:
                break
    flag = torch.tensor([1 if len(sizes) > 1 else 0])
    size, _ = get_comm_size_and_rank()
    if size > 1:
        import torch.distributed as dist
        flag = flag.to("cuda" if dist.get_backend() == "nccl" else "cpu")
        dist.all_reduce(flag, op=dist.ReduceOp.MAX)
    return bool(int(flag.item()))


def gather_deg(dataset):
    """In-degree histogram over the training set (PNA's `deg`), summed across ranks."""
    hist = torch.zeros(1, dtype=torch.long)
    for data in dataset:
        d = torch.bincount(data.edge_index[1], minlength=data.num_nodes)
        h = torch.bincount(d)
        if h.numel() > hist.numel():
            hist = torch.cat([hist, hist.new_zeros(h.numel() - hist.numel())])
        hist[:h.numel()] += h
    size, _ = get_comm_size_and_rank()
    if size > 1:
        import torch.distributed as dist
        dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
        n = torch.tensor([hist.numel()], device=dev)
        dist.all_reduce(n, op=dist.ReduceOp.MAX)
        padded = torch.zeros(int(n.item()), dtype=torch.long, device=dev)
        padded[:hist.numel()] = hist.to(dev)
        dist.all_reduce(padded, op=dist.ReduceOp.SUM)
        hist = padded.cpu()
    return hist


def update_config_edge_dim(arch):
    arch["edge_dim"] = None
    if arch.get("edge_features"):
        assert arch["mpnn_type"] in _EDGE_MODELS, \
            "Edge features can only be used with GAT, PNA, PNAPlus, PAINN, PNAEq, CGCNN, SchNet, EGNN, DimeNet, MACE."
        arch["edge_dim"] = len(arch["edge_features"])
    elif arch["mpnn_type"] == "CGCNN":
        arch["edge_dim"] = 0
    return arch


def update_config_equivariance(arch):
    if arch.get("equivariance"):
        assert arch["mpnn_type"] in _EQUIVARIANT_MODELS, \
            "E(3) equivariance can only be ensured for EGNN, SchNet, PNAEq, PAINN, and MACE."
    arch.setdefault("equivariance", False)
    return arch


def update_config_NN_outputs(nn_config, data, graph_size_variable):
    """Head dimensions from the first sample's y_loc; node heads divide by the node count."""
    kinds = nn_config["Variables_of_interest"]["type"]
    if nn_config["Training"]["compute_grad_energy"]:
        dims = nn_config["Variables_of_interest"]["output_dim"]
    elif getattr(data, "y_loc", None) is not None:
        dims = []
        for k, kind in enumerate(kinds):
            width = int(data.y_loc[0, k + 1]) - int(data.y_loc[0, k])
            if kind == "graph":
                dims.append(width)
            elif kind == "node":
                node_arch = nn_config["Architecture"]["output_heads"]["node"][0]["architecture"]
                if graph_size_variable and node_arch["type"] == "mlp_per_node":
                    raise ValueError('"mlp_per_node" is not allowed for variable graph size, Please set '
                                     'config["NeuralNetwork"]["Architecture"]["output_heads"]["node"]["type"] '
                                     'to be "mlp" or "conv" in input file.')
                dims.append(width // data.num_nodes)
            else:
                raise ValueError("Unknown output type", kind)
    else:
        if any(kind != "graph" for kind in kinds):
            raise ValueError("y_loc is needed for outputs that are not at graph levels")
        dims = nn_config["Variables_of_interest"]["output_dim"]
    arch = nn_config["Architecture"]
    arch["output_dim"], arch["output_type"], arch["num_nodes"] = dims, kinds, int(data.num_nodes)
    return nn_config


def update_config_minmax(dataset_path, var_config):
    """Min / max of the selected inputs and outputs (for de-normalisation of predictions)."""
    if "minmax_node_feature" in var_config or "minmax_graph_feature" in var_config:
        node_minmax, graph_minmax = var_config["minmax_node_feature"], var_config["minmax_graph_feature"]
    else:
        with open(dataset_path, "rb") as fh:
            node_minmax, graph_minmax = pickle.load(fh), pickle.load(fh)
    var_config["x_minmax"] = [node_minmax[:, i].tolist() for i in var_config["input_node_features"]]
    var_config["y_minmax"] = []
    for kind, index in zip(var_config["type"], var_config["output_index"]):
        if kind not in ("graph", "node"):
            raise ValueError("Unknown output type", kind)
        source = graph_minmax if kind == "graph" else node_minmax
        var_config["y_minmax"].append(source[:, index].tolist())
    return var_config


def _normalize_output_config(config):
    var = config["NeuralNetwork"]["Variables_of_interest"]
    if var.get("denormalize_output"):
        if var.get("minmax_node_feature") is not None and var.get("minmax_graph_feature") is not None:
            path = None
        else:
            first = list(config["Dataset"]["path"].values())[0]
            if first.endswith(".pkl"):
                path = first
            else:
                suffix = "" if "total" in config["Dataset"]["path"] else "_train"
                path = os.path.join(os.environ["SERIALIZED_DATA_PATH"], "serialized_dataset",
                                    config["Dataset"]["name"] + suffix + ".pkl")
        var = update_config_minmax(path, var)
    else:
        var["denormalize_output"] = False
    config["NeuralNetwork"]["Variables_of_interest"] = var
    return config


def update_config(config, train_loader, val_loader, test_loader):
    """Fill defaults and the fields derived from the data (input_dim, output_dim, pna_deg, edge_dim...)."""
    nn_config = config["NeuralNetwork"]
    arch, training = nn_config["Architecture"], nn_config["Training"]
    variable = _graph_size_variable(train_loader, val_loader, test_loader)
    first = _dataset_of(train_loader)[0]

    arch.setdefault("global_attn_engine", None)
    arch.setdefault("global_attn_type", None)
    arch.setdefault("global_attn_heads", 0)
    arch.setdefault("pe_dim", 0)
    arch["output_heads"] = update_multibranch_heads(arch["output_heads"])
    training.setdefault("compute_grad_energy", False)
    config["NeuralNetwork"] = update_config_NN_outputs(nn_config, first, variable)
    config = _normalize_output_config(config)

    arch["input_dim"] = len(nn_config["Variables_of_interest"]["input_node_features"])
    if arch["mpnn_type"] in ("PNA", "PNAPlus", "PNAEq"):
        dataset = _dataset_of(train_loader)
        deg = torch.as_tensor(dataset.pna_deg) if hasattr(dataset, "pna_deg") else gather_deg(dataset)
        arch["pna_deg"] = deg.tolist()
        arch["max_neighbours"] = len(arch["pna_deg"]) - 1
    else:
        arch["pna_deg"] = None
    arch["avg_num_neighbors"] = None
    for key in _ARCH_NONE_DEFAULTS:
        arch.setdefault(key, None)
    update_config_edge_dim(arch)
    update_config_equivariance(arch)
    for key, value in _ARCH_DEFAULTS.items():
        arch.setdefault(key, value)
    training.setdefault("conv_checkpointing", False)
    training.setdefault("loss_function_type", "mse")
    training.setdefault("Optimizer", {"type": "AdamW", "learning_rate": 1e-3})
    return config


def get_log_name_config(config):
    arch, training = config["NeuralNetwork"]["Architecture"], config["NeuralNetwork"]["Training"]
    name = config["Dataset"]["name"]
    cut = name.rfind("_")
    parts = [
        arch["mpnn_type"], "-r-", str(arch.get("radius")), "-ncl-", str(arch["num_conv_layers"]),
        "-hd-", str(arch["hidden_dim"]), "-ne-", str(training["num_epoch"]),
        "-lr-", str(training["Optimizer"]["learning_rate"]), "-bs-", str(training["batch_size"]),
        "-data-", name[:cut] if cut > 0 else name,
        "-node_ft-", "".join(str(i) for i in config["NeuralNetwork"]["Variables_of_interest"]["input_node_features"]),
        "-task_weights-", "".join(str(w) + "-" for w in arch["task_weights"]),
    ]
    return "".join(parts)


def save_config(config, log_name, path="./logs/"):
    if get_comm_size_and_rank()[1] == 0:
        os.makedirs(os.path.join(path, log_name), exist_ok=True)
        with open(os.path.join(path, log_name, "config.json"), "w") as fh:
            json.dump(config, fh, indent=4)


def merge_config(a: dict, b: dict) -> dict:
    out = deepcopy(a)
    for key, value in b.items():
        if isinstance(out.get(key), dict) and isinstance(value, dict):
            out[key] = merge_config(out[key], value)
        else:
            out[key] = deepcopy(value)
    return out
