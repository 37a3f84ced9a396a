"""Model factory with the reference's signatures (hydragnn/models/create.py:35-519).

`create_model_config(config["NeuralNetwork"], verbosity, use_gpu)` and the long positional
`create_model(...)` are kept verbatim so that existing drivers work; the four conv stacks of
the hot path (SchNet, EGNN, PNA, PAINN) are built on the sm_100a kernel library, every other
`mpnn_type` of the reference raises (they are outside this repo's scope).
"""
from typing import List, Union

import torch

from .EGCLStack import EGCLStack
from .PAINNStack import PAINNStack
from .PNAStack import PNAStack
from .PNAPlusStack import PNAPlusStack
from .PNAEqStack import PNAEqStack
from .SCFStack import SCFStack

# Architecture keys, in the positional order of create_model (create.py:40-81 of the reference)
_ARCH_KEYS = (
    "mpnn_type", "input_dim", "hidden_dim", "output_dim", "pe_dim", "global_attn_engine",
    "global_attn_type", "global_attn_heads", "output_type", "output_heads", "activation_function",
)
_ARCH_KEYS_TAIL = (
    "task_weights", "num_conv_layers", "freeze_conv_layers", "initial_bias", "num_nodes",
    "max_neighbours", "edge_dim", "pna_deg", "num_before_skip", "num_after_skip", "num_radial",
    "radial_type", "distance_transform", "basis_emb_size", "int_emb_size", "out_emb_size",
    "envelope_exponent", "num_spherical", "num_gaussians", "num_filters", "radius", "equivariance",
    "correlation", "max_ell", "node_max_ell", "avg_num_neighbors",
)
_NOT_ON_THIS_PATH = ("GIN", "GAT", "MFC", "SAGE", "CGCNN", "DimeNet", "MACE")


def create_model_config(config: dict, verbosity: int = 0, use_gpu: bool = True):
    arch, training = config["Architecture"], config["Training"]
    head = [arch[k] for k in _ARCH_KEYS]
    tail = [arch.get(k) for k in _ARCH_KEYS_TAIL]
    return create_model(*head, training["loss_function_type"], *tail,
                        training.get("conv_checkpointing", False), verbosity, use_gpu)


def _device(use_gpu):
    if use_gpu and torch.cuda.is_available():
        import os
        return torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)) % torch.cuda.device_count())
    return torch.device("cpu")


def create_model(
    mpnn_type: str,
    input_dim: int,
    hidden_dim: int,
    output_dim: list,
    pe_dim: int,
    global_attn_engine: str,
    global_attn_type: str,
    global_attn_heads: int,
    output_type: list,
    output_heads: dict,
    activation_function: str,
    loss_function_type: str,
    task_weights: list,
    num_conv_layers: int,
    freeze_conv: bool = False,
    initial_bias: float = None,
    num_nodes: int = None,
    max_neighbours: int = None,
    edge_dim: int = None,
    pna_deg: torch.tensor = None,
    num_before_skip: int = None,
    num_after_skip: int = None,
    num_radial: int = None,
    radial_type: str = None,
    distance_transform: str = None,
    basis_emb_size: int = None,
    int_emb_size: int = None,
    out_emb_size: int = None,
    envelope_exponent: int = None,
    num_spherical: int = None,
    num_gaussians: int = None,
    num_filters: int = None,
    radius: float = None,
    equivariance: bool = False,
    correlation: Union[int, List[int]] = None,
    max_ell: int = None,
    node_max_ell: int = None,
    avg_num_neighbors: int = None,
    conv_checkpointing: bool = False,
    verbosity: int = 0,
    use_gpu: bool = True,
):
    torch.manual_seed(0)   # same initial weights as the reference (create.py:131)
    shared = (input_dim, hidden_dim, output_dim, pe_dim, global_attn_engine, global_attn_type,
              global_attn_heads, output_type, output_heads, activation_function, loss_function_type,
              equivariance)
    options = dict(loss_weights=task_weights, freeze_conv=freeze_conv, initial_bias=initial_bias,
                   num_conv_layers=num_conv_layers, num_nodes=num_nodes)

    if mpnn_type == "PNA":
        assert pna_deg is not None, "PNA requires degree input."
        model = PNAStack("inv_node_feat, equiv_node_feat, edge_index", "inv_node_feat, edge_index",
                         pna_deg, edge_dim, *shared, **options)
    elif mpnn_type == "PNAPlus":
        assert pna_deg is not None, "PNAPlus requires degree input."
        assert envelope_exponent is not None, "PNAPlus requires envelope_exponent input."
        assert num_radial is not None, "PNAPlus requires num_radial input."
        assert radius is not None, "PNAPlus requires radius input."
        model = PNAPlusStack("inv_node_feat, equiv_node_feat, edge_index, rbf", "inv_node_feat, edge_index, rbf",
                             pna_deg, edge_dim, envelope_exponent, num_radial, radius, *shared, **options)
    elif mpnn_type == "SchNet":
        assert num_gaussians is not None, "SchNet requires num_guassians input."
        assert num_filters is not None, "SchNet requires num_filters input."
        assert radius is not None, "SchNet requires radius input."
        model = SCFStack("", "inv_node_feat, equiv_node_feat, edge_index, edge_weight, edge_rbf",
                         num_filters, edge_dim, num_gaussians, radius, *shared,
                         max_neighbours=max_neighbours, **options)
    elif mpnn_type == "EGNN":
        model = EGCLStack("inv_node_feat, equiv_node_feat, edge_index, edge_attr", "", edge_dim,
                          *shared, max_neighbours=max_neighbours, **options)
    elif mpnn_type == "PAINN":
        assert num_radial is not None, "PAINN requires num_radial input."
        assert radius is not None, "PAINN requires radius input."
        signature = "inv_node_feat, equiv_node_feat, edge_index, diff, dist"
        model = PAINNStack(signature, signature, edge_dim, num_radial, radius, *shared, **options)
    elif mpnn_type == "PNAEq":
        assert pna_deg is not None, "PNAEq requires degree input."
        signature = "inv_node_feat, equiv_node_feat, edge_index, edge_rbf, edge_vec"
        model = PNAEqStack(signature, signature, pna_deg, edge_dim, num_radial, radius, *shared, **options)
    elif mpnn_type in _NOT_ON_THIS_PATH:
        raise NotImplementedError(
            f"mpnn_type {mpnn_type} is not on the B200 hot path (SchNet, EGNN, PNA, PNAPlus, PNAEq, PAINN are)")
    else:
        raise ValueError("Unknown mpnn_type: {0}".format(mpnn_type))

    if conv_checkpointing:
        model.enable_conv_checkpointing()
    return model.to(_device(use_gpu))
