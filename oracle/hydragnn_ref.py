"""ORACLE (test infrastructure only -- never imported by the product path).

CPU restatement, in plain torch, of the reference's own hot-path modules, each
citing the reference file:line it follows.  It issues the same kind of ops the
reference issues on CPU (index_select, scatter_add_/index_add_, dense
nn.MultiheadAttention, nn.Linear) on top of oracle/pyg_lite.py, and keeps the
reference's module/parameter names so that state dicts are interchangeable
with both the reference and the product modules.

Pinning: the classes below are checked against golden vectors produced by the
REAL reference classes (imported from /root/reference with oracle/pyg_lite as
the stand-in for the absent third-party libraries) -- see tests/golden/
make_golden.py and tests/test_oracle_golden.py.  The third-party layer itself
is "parity unpinned" (see pyg_lite.py).
"""
import math
from typing import Optional

import numpy as np
import torch
from torch import nn

from . import pyg_lite as pyg


def _fdtype():
    """The reference writes `.float()` (Base.py:234,599-615).  Under the default dtype (fp32) this is
    the same cast; tests build an fp64 "truth" run of this oracle by switching the default dtype."""
    return torch.get_default_dtype()


# -- utils/model/operations.py:21-36 ------------------------------------------
def get_edge_vectors_and_lengths(positions, edge_index, shifts, normalize=False, eps=1e-9):
    vec = positions[edge_index[1]] - positions[edge_index[0]] + shifts
    length = torch.linalg.norm(vec, dim=-1, keepdim=True)
    if normalize:
        return vec / (length + eps), length
    return vec, length


# -- EGCLStack.py:292-298, utils/model/model.py:279-286 -------------------------
def unsorted_segment_sum(data, segment_ids, num_segments):
    out = data.new_zeros((num_segments, data.size(1)))
    return out.scatter_add_(0, segment_ids.unsqueeze(-1).expand(-1, data.size(1)), data)


def unsorted_segment_mean(data, segment_ids, num_segments):
    idx = segment_ids.unsqueeze(-1).expand(-1, data.size(1))
    total = data.new_zeros((num_segments, data.size(1))).scatter_add_(0, idx, data)
    count = data.new_zeros((num_segments, data.size(1))).scatter_add_(0, idx, torch.ones_like(data))
    return total / count.clamp(min=1)


# -- utils/model/model.py:29-62 ------------------------------------------------
def activation_function_selection(name):
    table = {
        "relu": nn.ReLU, "selu": nn.SELU, "prelu": nn.PReLU, "elu": nn.ELU,
        "lrelu_01": lambda: nn.LeakyReLU(0.1), "lrelu_025": lambda: nn.LeakyReLU(0.25),
        "lrelu_05": lambda: nn.LeakyReLU(0.5), "sigmoid": nn.Sigmoid,
    }
    return table[name]() if name in table else None


def loss_function_selection(name):
    if name == "mse":
        return torch.nn.functional.mse_loss
    if name == "mae":
        return torch.nn.functional.l1_loss
    if name == "rmse":
        return lambda x, y: torch.sqrt(torch.nn.functional.mse_loss(x, y))
    if name == "GaussianNLLLoss":
        return torch.nn.GaussianNLLLoss()
    raise ValueError(name)


# -- globalAtt/gps.py:32-152 ---------------------------------------------------
class GPSConv(nn.Module):
    def __init__(self, channels, conv, heads=1, dropout=0.0, act="relu", act_kwargs=None,
                 norm="batch_norm", norm_kwargs=None, attn_type="multihead", attn_kwargs=None):
        super().__init__()
        if attn_type != "multihead":
            raise ValueError(f"{attn_type} is not supported")
        self.channels, self.conv, self.heads, self.dropout = channels, conv, heads, dropout
        self.attn = nn.MultiheadAttention(channels, heads, batch_first=True, **(attn_kwargs or {}))
        self.mlp = nn.Sequential(
            nn.Linear(channels, 2 * channels), pyg.activation_resolver(act, **(act_kwargs or {})),
            nn.Dropout(dropout), nn.Linear(2 * channels, channels), nn.Dropout(dropout))
        self.norm1 = pyg.normalization_resolver(norm, channels, **(norm_kwargs or {}))
        self.norm2 = pyg.normalization_resolver(norm, channels, **(norm_kwargs or {}))
        self.norm3 = pyg.normalization_resolver(norm, channels, **(norm_kwargs or {}))

    def forward(self, inv_node_feat, equiv_node_feat, graph_batch=None, **kwargs):
        F = torch.nn.functional
        branches = []
        if self.conv is not None:                                           # :112-123
            h, equiv_node_feat = self.conv(inv_node_feat=inv_node_feat,
                                           equiv_node_feat=equiv_node_feat, **kwargs)
            h = F.dropout(h, p=self.dropout, training=self.training) + inv_node_feat
            branches.append(self.norm1(h) if self.norm1 is not None else h)
        dense, mask = pyg.to_dense_batch(inv_node_feat, graph_batch)        # :126
        h, _ = self.attn(dense, dense, dense, key_padding_mask=~mask, need_weights=False)
        h = F.dropout(h[mask], p=self.dropout, training=self.training) + inv_node_feat
        branches.append(self.norm2(h) if self.norm2 is not None else h)
        out = sum(branches)                                                 # :143
        out = out + self.mlp(out)
        if self.norm3 is not None:
            out = self.norm3(out)
        return out, equiv_node_feat


# -- models/Base.py:692-752 ----------------------------------------------------
class MLPNode(nn.Module):
    def __init__(self, input_dim, output_dim, num_mlp, hidden_dim_node, node_type, activation_function):
        super().__init__()
        self.input_dim, self.output_dim, self.node_type = input_dim, output_dim, node_type
        self.num_mlp = self.num_nodes = num_mlp
        self.mlp = nn.ModuleList()
        for _ in range(num_mlp):
            dims = [input_dim] + list(hidden_dim_node)
            layers = []
            for a, b in zip(dims[:-1], dims[1:]):
                layers += [nn.Linear(a, b), activation_function]
            layers.append(nn.Linear(dims[-1], output_dim))
            self.mlp.append(nn.Sequential(*layers))

    def forward(self, x, batch):
        if self.node_type == "mlp":
            return self.mlp[0](x)
        outs = x.new_zeros((x.shape[0], self.output_dim))
        for inode in range(self.num_nodes):
            rows = torch.arange(inode, batch.shape[0], self.num_nodes)
            outs[rows] = self.mlp[inode](x[rows])
        return outs


# -- models/Base.py:31-686 -----------------------------------------------------
class Base(nn.Module):
    def __init__(self, input_args, conv_args, input_dim, hidden_dim, output_dim, pe_dim,
                 global_attn_engine, global_attn_type, global_attn_heads, output_type, config_heads,
                 activation_function_type, loss_function_type, equivariance, ilossweights_hyperp=1,
                 loss_weights=(1.0, 1.0, 1.0), ilossweights_nll=0, freeze_conv=False,
                 initial_bias=None, dropout=0.25, num_conv_layers=16, num_nodes=None):
        super().__init__()
        self.input_args, self.conv_args = input_args, conv_args
        self.global_attn_engine, self.global_attn_type = global_attn_engine, global_attn_type
        self.global_attn_heads, self.global_attn_dropout = global_attn_heads, dropout
        self.input_dim, self.pe_dim, self.hidden_dim = input_dim, pe_dim, hidden_dim
        self.dropout, self.num_conv_layers, self.num_nodes = dropout, num_conv_layers, num_nodes
        self.graph_convs, self.feature_layers, self.heads_NN = nn.ModuleList(), nn.ModuleList(), nn.ModuleList()
        self.config_heads, self.head_type, self.head_dims = config_heads, output_type, output_dim
        self.num_heads = len(output_dim)
        self.convs_node_hidden, self.batch_norms_node_hidden = nn.ModuleDict(), nn.ModuleDict()
        self.convs_node_output, self.batch_norms_node_output = nn.ModuleDict(), nn.ModuleDict()
        self.equivariance = equivariance
        self.activation_function = activation_function_selection(activation_function_type)
        self.var_output = 1 if loss_function_type == "GaussianNLLLoss" else 0
        self.loss_function_type = loss_function_type
        self.loss_function = loss_function_selection(loss_function_type)
        self.ilossweights_nll, self.ilossweights_hyperp = ilossweights_nll, ilossweights_hyperp
        if ilossweights_hyperp * ilossweights_nll == 1:
            raise ValueError("ilossweights_hyperp and ilossweights_nll cannot be both set to 1.")
        if ilossweights_hyperp == 1:
            if len(loss_weights) != self.num_heads:
                raise ValueError("Inconsistent number of loss weights and tasks")
            total = sum(abs(w) for w in loss_weights)
            self.loss_weights = [w / total for w in loss_weights]              # :113-114
        self.use_edge_attr = False                                              # :116-127
        if getattr(self, "edge_dim", None) is not None and self.edge_dim > 0:
            self.use_edge_attr = True
            if "edge_attr" not in self.input_args:
                self.input_args += ", edge_attr"
            if "edge_attr" not in self.conv_args:
                self.conv_args += ", edge_attr"
        self.freeze_conv, self.initial_bias = freeze_conv, initial_bias
        if self.global_attn_engine:                                             # :136-154
            self.use_global_attn = True
            self.embed_dim = self.edge_embed_dim = hidden_dim
            if self.is_edge_model:
                if "edge_attr" not in self.input_args:
                    self.input_args += ", edge_attr"
                if "edge_attr" not in self.conv_args:
                    self.conv_args += ", edge_attr"
        else:
            self.use_global_attn = False
            self.embed_dim = input_dim
            self.edge_embed_dim = getattr(self, "edge_dim", None)
        if self.use_global_attn:                                                # :161-172
            self.pos_emb = nn.Linear(pe_dim, hidden_dim, bias=False)
            if input_dim:
                self.node_emb = nn.Linear(input_dim, hidden_dim, bias=False)
                self.node_lin = nn.Linear(2 * hidden_dim, hidden_dim, bias=False)
            if self.is_edge_model:
                self.rel_pos_emb = nn.Linear(pe_dim, hidden_dim, bias=False)
                if self.use_edge_attr:
                    self.edge_emb = nn.Linear(self.edge_dim, hidden_dim, bias=False)
                    self.edge_lin = nn.Linear(2 * hidden_dim, hidden_dim, bias=False)
        self._init_conv()
        if freeze_conv:
            for group in (self.graph_convs, self.feature_layers):
                for p in group.parameters():
                    p.requires_grad = False
        self._multihead()
        if initial_bias is not None:
            for head, kind in zip(self.heads_NN, self.head_type):
                if kind == "graph":
                    head["branch-0"][-1].bias.data.fill_(initial_bias)

    # :183-196
    def _apply_global_attn(self, mpnn):
        if self.use_global_attn and self.global_attn_engine == "GPS":
            return GPSConv(channels=self.hidden_dim, conv=mpnn, heads=self.global_attn_heads,
                           dropout=self.global_attn_dropout, attn_type=self.global_attn_type)
        return mpnn

    # :198-215
    def _init_conv(self):
        dims = [self.embed_dim] + [self.hidden_dim] * self.num_conv_layers
        for a, b in zip(dims[:-1], dims[1:]):
            self.graph_convs.append(self._apply_global_attn(self.get_conv(a, b, edge_dim=self.edge_embed_dim)))
            self.feature_layers.append(pyg.BatchNorm(self.hidden_dim))

    def _gps_embed(self, data, conv_args):                                     # :229-243
        x = self.pos_emb(data.pe)
        if self.input_dim:
            x = self.node_lin(torch.cat((self.node_emb(data.x.to(_fdtype())), x), 1))
        if self.is_edge_model:
            e = self.rel_pos_emb(data.rel_pe)
            if self.use_edge_attr:
                e = self.edge_lin(torch.cat((self.edge_emb(conv_args["edge_attr"]), e), 1))
            conv_args["edge_attr"] = e
        return x

    def _embedding(self, data):                                                # :217-245
        if not hasattr(data, "edge_shifts") or data.edge_shifts is None:
            data.edge_shifts = torch.zeros((data.edge_index.size(1), 3))
        conv_args = {"edge_index": data.edge_index.to(torch.long)}
        if self.use_edge_attr:
            conv_args["edge_attr"] = data.edge_attr
        if self.use_global_attn:
            return self._gps_embed(data, conv_args), data.pos, conv_args
        return data.x, data.pos, conv_args

    def _conv_accepts_last_layer(self):
        import inspect
        return "last_layer" in inspect.signature(self.get_conv).parameters

    def _node_conv(self, a, b, last):
        return self.get_conv(a, b, last_layer=last) if self._conv_accepts_last_layer() else self.get_conv(a, b)

    def _init_node_conv(self):                                                 # :260-340
        branches = self.config_heads["node"]
        if any(b["architecture"]["type"] != "conv" for b in branches):
            return
        node_heads = [i for i, t in enumerate(self.head_type) if t == "node"]
        if not node_heads:
            return
        for b in branches:
            arch, name = b["architecture"], b["type"]
            dims = arch["dim_headlayers"]
            hidden, hidden_bn, outs, outs_bn = nn.ModuleList(), nn.ModuleList(), nn.ModuleList(), nn.ModuleList()
            hidden.append(self._node_conv(self.hidden_dim, dims[0], False))
            hidden_bn.append(pyg.BatchNorm(dims[0]))
            for k in range(arch["num_headlayers"] - 1):
                hidden.append(self._node_conv(dims[k], dims[k + 1], False))
                hidden_bn.append(pyg.BatchNorm(dims[k + 1]))
            for ihead in node_heads:
                width = self.head_dims[ihead] * (1 + self.var_output)
                outs.append(self._node_conv(dims[-1], width, True))
                outs_bn.append(pyg.BatchNorm(width))
            self.convs_node_hidden[name], self.batch_norms_node_hidden[name] = hidden, hidden_bn
            self.convs_node_output[name], self.batch_norms_node_output[name] = outs, outs_bn

    def _multihead(self):                                                      # :342-440
        self.graph_shared = nn.ModuleDict()
        self.num_branches = 1
        if "graph" in self.config_heads:
            self.num_branches = len(self.config_heads["graph"])
            for b in self.config_heads["graph"]:
                arch = b["architecture"]
                width = arch["dim_sharedlayers"]
                layers = [nn.Linear(self.hidden_dim, width), self.activation_function]
                for _ in range(arch["num_sharedlayers"] - 1):
                    layers += [nn.Linear(width, width), self.activation_function]
                self.graph_shared[b["type"]] = nn.Sequential(*layers)
        if "node" in self.config_heads:
            self._init_node_conv()
        inode = 0
        for ihead in range(self.num_heads):
            head = nn.ModuleDict()
            width_out = self.head_dims[ihead] * (1 + self.var_output)
            if self.head_type[ihead] == "graph":
                for b in self.config_heads["graph"]:
                    arch = b["architecture"]
                    dh = arch["dim_headlayers"]
                    layers = [nn.Linear(arch["dim_sharedlayers"], dh[0]), self.activation_function]
                    for k in range(arch["num_headlayers"] - 1):
                        layers += [nn.Linear(dh[k], dh[k + 1]), self.activation_function]
                    layers.append(nn.Linear(dh[-1], width_out))
                    head[b["type"]] = nn.Sequential(*layers)
            elif self.head_type[ihead] == "node":
                for b in self.config_heads["node"]:
                    arch, name = b["architecture"], b["type"]
                    kind = arch["type"]
                    if kind in ("mlp", "mlp_per_node"):
                        self.num_mlp = 1 if kind == "mlp" else self.num_nodes
                        assert self.num_nodes is not None, "num_nodes must be positive integer for MLP"
                        head[name] = MLPNode(self.hidden_dim, width_out, self.num_mlp,
                                             arch["dim_headlayers"], kind, self.activation_function)
                    elif kind == "conv":
                        seq = nn.ModuleList()
                        for conv, bn in zip(self.convs_node_hidden[name], self.batch_norms_node_hidden[name]):
                            seq.append(conv)
                            seq.append(bn)
                        seq.append(self.convs_node_output[name][inode])
                        seq.append(self.batch_norms_node_output[name][inode])
                        head[name] = seq
                        inode += 1
                    else:
                        raise ValueError("Unknown head NN structure for node features" + kind)
            else:
                raise ValueError("Unknown head type" + self.head_type[ihead])
            self.heads_NN.append(head)

    def forward(self, data):                                                   # :446-570
        x, equiv, conv_args = self._embedding(data)
        for conv, feat in zip(self.graph_convs, self.feature_layers):
            x, equiv = conv(inv_node_feat=x, equiv_node_feat=equiv, **conv_args)
            x = self.activation_function(feat(x))
        if data.batch is None:
            x_graph = x.mean(dim=0, keepdim=True)
            data.batch = torch.zeros(x.size(0), dtype=torch.long)
        else:
            x_graph = pyg.global_mean_pool(x, data.batch)
        if self.num_branches != 1:
            raise NotImplementedError("oracle restates single-branch heads only")
        outputs, outputs_var = [], []
        for head_dim, head, kind in zip(self.head_dims, self.heads_NN, self.head_type):
            if kind == "graph":
                out = head["branch-0"](self.graph_shared["branch-0"](x_graph))
            else:
                node_kind = self.config_heads["node"][0]["architecture"]["type"]
                if node_kind == "conv":
                    h, eq = x, equiv
                    for conv, bn in zip(head["branch-0"][0::2], head["branch-0"][1::2]):
                        h, eq = conv(inv_node_feat=h, equiv_node_feat=eq, **conv_args)
                        h = self.activation_function(bn(h))
                    out = h
                else:
                    out = head["branch-0"](x=x, batch=data.batch)
            outputs.append(out[:, :head_dim])
            outputs_var.append(out[:, head_dim:] ** 2)
        return (outputs, outputs_var) if self.var_output else outputs

    def loss(self, pred, value, head_index):                                   # :572-580
        var = None
        if self.var_output:
            pred, var = pred
        return self.loss_hpweighted(pred, value, head_index, var=var)

    def loss_hpweighted(self, pred, value, head_index, var=None):              # :659-686
        tot, tasks = 0, []
        for ihead in range(self.num_heads):
            p = pred[ihead]
            v = value[head_index[ihead]]
            if p.shape != v.shape:
                v = torch.reshape(v, p.shape)
            term = self.loss_function(p, v) if var is None else self.loss_function(p, v, var[ihead])
            tot = tot + term * self.loss_weights[ihead]
            tasks.append(term)
        return tot, tasks

    def energy_force_loss(self, pred, data):                                   # :582-636
        assert data.pos.requires_grad
        assert self.num_heads == 1 and self.head_type[0] == "node"
        e_graph = pyg.scatter_add(pred[0], data.batch, dim=0).squeeze().to(_fdtype())
        e_true = data.energy.squeeze().to(_fdtype())
        w_e = self.loss_weights[0]
        e_loss = self.loss_function(e_graph, e_true)
        f_true = data.forces.to(_fdtype())
        f_pred = -torch.autograd.grad(e_graph, data.pos, grad_outputs=torch.ones_like(e_graph),
                                      retain_graph=e_graph.requires_grad, create_graph=True)[0].to(_fdtype())
        w_f = w_e * torch.mean(torch.abs(e_true)) / (torch.mean(torch.abs(f_true)) + 1e-8)
        tot = e_loss * w_e + self.loss_function(f_pred, f_true) * w_f
        self.last_forces_pred = f_pred          # kept for the parity tests
        return tot, [e_loss]


# -- models/SCFStack.py ---------------------------------------------------------
class CFConv(pyg.MessagePassing):                                              # :214-293
    def __init__(self, in_channels, out_channels, num_filters, nn_, cutoff, equivariant):
        super().__init__(aggr="add")
        self.lin1 = nn.Linear(in_channels, num_filters, bias=False)
        self.lin2 = nn.Linear(num_filters, out_channels)
        self.nn, self.cutoff, self.equivariant = nn_, cutoff, equivariant
        if equivariant:
            last = nn.Linear(num_filters, 1, bias=False)
            nn.init.xavier_uniform_(last.weight, gain=0.001)
            self.coord_mlp = nn.Sequential(nn.Linear(num_filters, num_filters), nn.ReLU(), last)
        nn.init.xavier_uniform_(self.lin1.weight)
        nn.init.xavier_uniform_(self.lin2.weight)
        self.lin2.bias.data.fill_(0)

    def forward(self, x, pos, edge_index, edge_weight, edge_rbf, edge_attr=None):
        C = 0.5 * (torch.cos(edge_weight * math.pi / self.cutoff) + 1.0)     # :268 (no d<cutoff mask)
        feats = edge_rbf if edge_attr is None else torch.cat([edge_rbf, edge_attr], dim=-1)
        W = self.nn(feats) * C.view(-1, 1)
        x = self.lin1(x)
        if self.equivariant:                                                   # :276-283, 244-252
            shifts = torch.zeros((edge_index.size(1), 3))
            unit, _ = get_edge_vectors_and_lengths(pos, edge_index, shifts, normalize=True, eps=1.0)
            trans = torch.clamp(unit * self.coord_mlp(W), min=-100, max=100)
            pos = pos + unsorted_segment_mean(trans, edge_index[0], num_segments=pos.size(0))
        x = self.lin2(self.propagate(edge_index, x=x, W=W))
        return (x, pos) if self.equivariant else x

    def message(self, x_j, W):
        return x_j * W


class SCFStack(Base):                                                          # :34-211
    def __init__(self, input_args, conv_args, num_filters, edge_dim, num_gaussians, radius, *args,
                 max_neighbours: Optional[int] = None, **kwargs):
        self.radius, self.max_neighbours = radius, max_neighbours
        self.num_filters, self.edge_dim, self.num_gaussians = num_filters, edge_dim, num_gaussians
        self.is_edge_model = True
        super().__init__(input_args, conv_args, *args, **kwargs)

    def _init_conv(self):
        self.distance_expansion = pyg.GaussianSmearing(0.0, self.radius, self.num_gaussians)
        self.interaction_graph = pyg.RadiusInteractionGraph(self.radius, self.max_neighbours)
        dims = [self.embed_dim] + [self.hidden_dim] * self.num_conv_layers
        for k, (a, b) in enumerate(zip(dims[:-1], dims[1:])):
            last = k == self.num_conv_layers - 1
            self.graph_convs.append(self._apply_global_attn(self.get_conv(a, b, last, edge_dim=self.edge_embed_dim)))
            self.feature_layers.append(nn.Identity())

    def get_conv(self, input_dim, output_dim, last_layer, edge_dim=None):      # :89-153
        width = self.num_gaussians + edge_dim if edge_dim else self.num_gaussians
        mlp = nn.Sequential(nn.Linear(width, self.num_filters), pyg.ShiftedSoftplus(),
                            nn.Linear(self.num_filters, self.num_filters))
        inter = CFConv(input_dim, output_dim, self.num_filters, mlp, self.radius,
                       self.equivariance and not last_layer)
        passthrough = (lambda inv_node_feat, equiv_node_feat: [inv_node_feat, equiv_node_feat],
                       "inv_node_feat, equiv_node_feat -> inv_node_feat, equiv_node_feat")
        if self.use_edge_attr or (self.use_global_attn and self.is_edge_model):
            return pyg.Sequential(
                "inv_node_feat, equiv_node_feat, edge_index, edge_weight, edge_rbf, batch, edge_attr",
                [(inter, self.conv_args + " -> inv_node_feat"), passthrough])
        graph_steps = [(self.interaction_graph, "equiv_node_feat, batch -> edge_index, edge_weight"),
                       (self.distance_expansion, "edge_weight -> edge_rbf")]
        if self.equivariance and not last_layer:
            return pyg.Sequential("inv_node_feat, equiv_node_feat, batch", graph_steps + [
                (inter, self.conv_args + " -> inv_node_feat, equiv_node_feat")])
        return pyg.Sequential("inv_node_feat, equiv_node_feat, batch", graph_steps + [
            (inter, self.conv_args + " -> inv_node_feat"), passthrough])

    def _embedding(self, data):                                                # :155-208
        Base._embedding(self, data)
        edge_mode = self.use_edge_attr or (self.use_global_attn and self.is_edge_model)
        if edge_mode and self.equivariance:
            raise Exception("SchNet: edge attributes / GPS encodings are incompatible with equivariance")
        if edge_mode:
            data.edge_shifts = torch.zeros((data.edge_index.size(1), 3))       # :166-168 (PBC dropped)
            _, dist = get_edge_vectors_and_lengths(data.pos, data.edge_index, data.edge_shifts)
            w = dist.squeeze()
            conv_args = {"edge_index": data.edge_index, "edge_weight": w,
                         "edge_rbf": self.distance_expansion(w), "batch": data.batch}
            if self.use_edge_attr:
                conv_args["edge_attr"] = data.edge_attr
        else:
            conv_args = {"batch": data.batch}
        if self.use_global_attn:
            return self._gps_embed(data, conv_args), data.pos, conv_args
        return data.x, data.pos, conv_args


# -- models/EGCLStack.py --------------------------------------------------------
class E_GCL(nn.Module):                                                        # :175-289
    def __init__(self, input_channels, output_channels, hidden_channels, edge_attr_dim=0,
                 equivariant=False):
        super().__init__()
        act = nn.ReLU()
        self.equivariant, self.edge_attr_dim = equivariant, edge_attr_dim
        self.edge_mlp = nn.Sequential(
            nn.Linear(2 * input_channels + 1 + edge_attr_dim, hidden_channels), act,
            nn.Linear(hidden_channels, hidden_channels), act)
        self.node_mlp = nn.Sequential(
            nn.Linear(hidden_channels + input_channels, hidden_channels), act,
            nn.Linear(hidden_channels, output_channels))
        if equivariant:
            last = nn.Linear(hidden_channels, 1, bias=False)
            nn.init.xavier_uniform_(last.weight, gain=0.001)
            self.coord_mlp = nn.Sequential(nn.Linear(hidden_channels, hidden_channels), act, last, nn.Tanh())

    def forward(self, x, coord, edge_index, edge_attr, node_attr=None):
        row, col = edge_index
        shifts = torch.zeros((edge_index.size(1), 3))                          # :275-277
        unit, radial = get_edge_vectors_and_lengths(coord, edge_index, shifts, normalize=True, eps=1.0)
        parts = [x[row], x[col], radial] + ([] if edge_attr is None else [edge_attr])
        m = self.edge_mlp(torch.cat(parts, dim=1))                             # :240-245
        if self.equivariant:                                                   # :263-271
            trans = torch.clamp(unit * self.coord_mlp(m), min=-100, max=100)
            coord = coord + unsorted_segment_mean(trans, row, num_segments=coord.size(0)) * 1.0
        agg = unsorted_segment_sum(m, row, num_segments=x.size(0))             # :253
        out = self.node_mlp(torch.cat([x, agg], dim=1))
        return (out, coord) if self.equivariant else out


class EGCLStack(Base):                                                         # :22-147
    def __init__(self, input_args, conv_args, edge_attr_dim, *args, max_neighbours=None, **kwargs):
        self.edge_dim = 0 if edge_attr_dim is None else edge_attr_dim
        self.is_edge_model = True
        super().__init__(input_args, conv_args, *args, **kwargs)
        assert self.input_args == "inv_node_feat, equiv_node_feat, edge_index, edge_attr"

    def _init_conv(self):
        dims = [self.embed_dim] + [self.hidden_dim] * self.num_conv_layers
        for k, (a, b) in enumerate(zip(dims[:-1], dims[1:])):
            last = k == self.num_conv_layers - 1
            self.graph_convs.append(self._apply_global_attn(self.get_conv(a, b, last, edge_dim=self.edge_embed_dim)))
            self.feature_layers.append(nn.Identity())

    def get_conv(self, input_dim, output_dim, last_layer=False, edge_dim=None):
        if not edge_dim:
            edge_dim = self.edge_dim
        eq = self.equivariance and not last_layer
        egcl = E_GCL(input_dim, output_dim, self.hidden_dim, edge_attr_dim=edge_dim, equivariant=eq)
        sig = "inv_node_feat, equiv_node_feat, edge_index, edge_attr"
        if eq:
            return pyg.Sequential(self.input_args, [(egcl, sig + " -> inv_node_feat, equiv_node_feat")])
        return pyg.Sequential(self.input_args, [
            (egcl, sig + " -> inv_node_feat"),
            (lambda inv_node_feat, equiv_node_feat: [inv_node_feat, equiv_node_feat],
             "inv_node_feat, equiv_node_feat -> inv_node_feat, equiv_node_feat")])

    def _embedding(self, data):                                                # :110-144
        Base._embedding(self, data)
        data.edge_shifts = torch.zeros((data.edge_index.size(1), 3))
        conv_args = {"edge_index": data.edge_index,
                     "edge_attr": data.edge_attr if self.edge_dim > 0 else None}
        if self.use_global_attn:
            return self._gps_embed(data, conv_args), data.pos, conv_args
        return data.x, data.pos, conv_args


# -- models/PNAStack.py ---------------------------------------------------------
class PNAStack(Base):                                                          # :19-70
    def __init__(self, input_args, conv_args, deg, edge_dim, *args, **kwargs):
        self.aggregators = ["mean", "min", "max", "std"]
        self.scalers = ["identity", "amplification", "attenuation", "linear"]
        self.deg = torch.Tensor(deg)
        self.edge_dim = edge_dim
        self.is_edge_model = True
        super().__init__(input_args, conv_args, *args, **kwargs)

    def get_conv(self, input_dim, output_dim, edge_dim=None):
        pna = pyg.PNAConv(in_channels=input_dim, out_channels=output_dim, aggregators=self.aggregators,
                          scalers=self.scalers, deg=self.deg, edge_dim=edge_dim, pre_layers=1,
                          post_layers=1, divide_input=False)
        return pyg.Sequential(self.input_args, [
            (pna, self.conv_args + " -> inv_node_feat"),
            (lambda inv_node_feat, equiv_node_feat: [inv_node_feat, equiv_node_feat],
             "inv_node_feat, equiv_node_feat -> inv_node_feat, equiv_node_feat")])


# -- models/PNAPlusStack.py -----------------------------------------------------
class PNAPlusConv(pyg.MessagePassing):                                         # :144-283
    """PNAConv with a radial basis: messages pre([x_i, x_j, e]) * rbf_lin(rbf), e = edge_encoder(
    [edge_attr, rbf_emb(rbf)]) or rbf_emb(rbf); towers = 1, one pre / post Linear."""

    def __init__(self, in_channels, out_channels, aggregators, scalers, deg, edge_dim=None, num_radial=5):
        super().__init__(aggr=pyg.DegreeScalerAggregation(aggregators, scalers, deg), node_dim=0)
        self.in_channels, self.out_channels, self.edge_dim = in_channels, out_channels, edge_dim
        F_in = in_channels
        self.pre_nns = nn.ModuleList([nn.Sequential(nn.Linear(3 * F_in, F_in))])
        self.post_nns = nn.ModuleList([nn.Sequential(
            nn.Linear((len(aggregators) * len(scalers) + 1) * F_in, out_channels))])
        self.lin = nn.Linear(out_channels, out_channels)
        self.rbf_lin = nn.Linear(num_radial, F_in, bias=False)
        self.rbf_emb = nn.Sequential(nn.Linear(num_radial, F_in), nn.ReLU())
        if edge_dim is not None:
            self.edge_encoder = nn.Linear(F_in + edge_dim, F_in)

    def forward(self, x, edge_index, rbf=None, edge_attr=None):                # :239-256
        x = x.view(-1, 1, self.in_channels)
        out = self.propagate(edge_index, x=x, edge_attr=edge_attr, rbf=rbf)
        out = torch.cat([x, out], dim=-1)
        return self.lin(self.post_nns[0](out[:, 0]))

    def message(self, x_i, x_j, rbf=None, edge_attr=None):                     # :258-283
        e = self.rbf_emb(rbf)
        if edge_attr is not None:
            e = self.edge_encoder(torch.cat([edge_attr, e], dim=-1))
        h = torch.cat([x_i, x_j, e.view(-1, 1, self.in_channels)], dim=-1)
        return self.pre_nns[0](h[:, 0]).unsqueeze(1) * self.rbf_lin(rbf).view(-1, 1, self.in_channels)


class PNAPlusStack(Base):                                                      # :40-141
    def __init__(self, input_args, conv_args, deg, edge_dim, envelope_exponent, num_radial, radius,
                 *args, **kwargs):
        self.aggregators = ["mean", "min", "max", "std"]
        self.scalers = ["identity", "amplification", "attenuation", "linear"]
        self.deg = torch.Tensor(deg)
        self.edge_dim, self.envelope_exponent = edge_dim, envelope_exponent
        self.num_radial, self.radius = num_radial, radius
        self.is_edge_model = True
        super().__init__(input_args, conv_args, *args, **kwargs)
        self.rbf = pyg.BesselBasisLayer(num_radial, radius, envelope_exponent)

    def get_conv(self, input_dim, output_dim, edge_dim=None):
        pna = PNAPlusConv(input_dim, output_dim, self.aggregators, self.scalers, self.deg, edge_dim,
                          self.num_radial)
        return pyg.Sequential(self.input_args, [
            (pna, self.conv_args + " -> inv_node_feat"),
            (lambda inv_node_feat, equiv_node_feat: [inv_node_feat, equiv_node_feat],
             "inv_node_feat, equiv_node_feat -> inv_node_feat, equiv_node_feat")])

    def _embedding(self, data):                                                # :104-138
        Base._embedding(self, data)
        _, dist = get_edge_vectors_and_lengths(data.pos, data.edge_index, data.edge_shifts)
        conv_args = {"edge_index": data.edge_index.to(torch.long), "rbf": self.rbf(dist.squeeze())}
        if self.use_edge_attr:
            conv_args["edge_attr"] = data.edge_attr
        if self.use_global_attn:
            return self._gps_embed(data, conv_args), data.pos, conv_args
        return data.x, data.pos, conv_args


# -- models/PAINNStack.py -------------------------------------------------------
def sinc_expansion(edge_dist, num_radial, cutoff):                             # :322-329
    n = torch.arange(num_radial, device=edge_dist.device) + 1
    return torch.sin(edge_dist * n * torch.pi / cutoff) / edge_dist


def cosine_cutoff(edge_dist, cutoff):                                          # :332-343
    return torch.where(edge_dist < cutoff, 0.5 * (torch.cos(torch.pi * edge_dist / cutoff) + 1),
                       torch.tensor(0.0, dtype=edge_dist.dtype))


class PainnMessage(nn.Module):                                                 # :194-263
    def __init__(self, node_size, num_radial, cutoff, edge_dim):
        super().__init__()
        self.node_size, self.num_radial, self.cutoff, self.edge_dim = node_size, num_radial, cutoff, edge_dim
        self.scalar_message_mlp = nn.Sequential(nn.Linear(node_size, node_size), nn.SiLU(),
                                                nn.Linear(node_size, node_size * 3))
        self.filter_layer = nn.Linear(num_radial, node_size * 3)
        if edge_dim is not None:
            self.edge_filter = nn.Sequential(nn.Linear(edge_dim, node_size), nn.SiLU(),
                                             nn.Linear(node_size, node_size * 3))

    def forward(self, node_scalar, node_vector, edge, edge_diff, edge_dist, edge_attr=None):
        fw = self.filter_layer(sinc_expansion(edge_dist, self.num_radial, self.cutoff))
        fw = fw * cosine_cutoff(edge_dist, self.cutoff)
        if edge_attr is not None:
            fw = fw * self.edge_filter(edge_attr)
        fo = fw * self.scalar_message_mlp(node_scalar)[edge[:, 1]]
        g_sv, g_ev, m_s = torch.split(fo, self.node_size, dim=1)
        m_v = node_vector[edge[:, 1]] * g_sv.unsqueeze(1)
        m_v = m_v + g_ev.unsqueeze(1) * (edge_diff / edge_dist).unsqueeze(-1)  # :247-251 (second /d)
        ds = torch.zeros_like(node_scalar).index_add_(0, edge[:, 0], m_s)
        dv = torch.zeros_like(node_vector).index_add_(0, edge[:, 0], m_v)
        return node_scalar + ds, node_vector + dv


class PainnUpdate(nn.Module):                                                  # :266-319
    def __init__(self, node_size, last_layer=False):
        super().__init__()
        self.update_U = nn.Linear(node_size, node_size)
        self.update_V = nn.Linear(node_size, node_size)
        self.last_layer = last_layer
        self.update_mlp = nn.Sequential(nn.Linear(node_size * 2, node_size), nn.SiLU(),
                                        nn.Linear(node_size, node_size * (2 if last_layer else 3)))

    def forward(self, node_scalar, node_vector):
        Uv, Vv = self.update_U(node_vector), self.update_V(node_vector)
        a = self.update_mlp(torch.cat((torch.linalg.norm(Vv, dim=1), node_scalar), dim=1))
        inner = torch.sum(Uv * Vv, dim=1)
        if self.last_layer:
            a_sv, a_ss = torch.split(a, node_vector.shape[-1], dim=1)
            return node_scalar + a_sv * inner + a_ss
        a_vv, a_sv, a_ss = torch.split(a, node_vector.shape[-1], dim=1)
        return node_scalar + a_sv * inner + a_ss, node_vector + a_vv.unsqueeze(1) * Uv


class PAINNStack(Base):                                                        # :27-191
    def __init__(self, input_args, conv_args, edge_dim, num_radial, radius, *args, **kwargs):
        self.edge_dim, self.num_radial, self.radius = edge_dim, num_radial, radius
        self.is_edge_model = True
        super().__init__(input_args, conv_args, *args, **kwargs)

    def _init_conv(self):
        dims = [self.embed_dim] + [self.hidden_dim] * self.num_conv_layers
        for k, (a, b) in enumerate(zip(dims[:-1], dims[1:])):
            last = k == self.num_conv_layers - 1
            self.graph_convs.append(self._apply_global_attn(self.get_conv(a, b, last, edge_dim=self.edge_embed_dim)))
            self.feature_layers.append(nn.Identity())

    def get_conv(self, input_dim, output_dim, last_layer=False, edge_dim=None):  # :76-147
        assert (output_dim if input_dim == 1 else input_dim) > 1
        msg = PainnMessage(input_dim, self.num_radial, self.radius, edge_dim)
        upd = PainnUpdate(input_dim, last_layer)
        s_out = nn.Sequential(nn.Linear(input_dim, output_dim), nn.Tanh(), nn.Linear(output_dim, output_dim))
        both = "inv_node_feat, equiv_node_feat -> inv_node_feat, equiv_node_feat"
        steps = [(msg, self.conv_args + " -> inv_node_feat, equiv_node_feat")]
        if not last_layer:
            steps += [(upd, both), (s_out, "inv_node_feat -> inv_node_feat"),
                      (nn.Linear(input_dim, output_dim), "equiv_node_feat -> equiv_node_feat")]
        else:
            steps += [(upd, "inv_node_feat, equiv_node_feat -> inv_node_feat"),
                      (s_out, "inv_node_feat -> inv_node_feat")]
        steps.append((lambda inv_node_feat, equiv_node_feat: [inv_node_feat, equiv_node_feat], both))
        return pyg.Sequential(self.input_args, steps)

    def _embedding(self, data):                                                # :149-191
        Base._embedding(self, data)
        unit, dist = get_edge_vectors_and_lengths(data.pos, data.edge_index, data.edge_shifts, normalize=True)
        conv_args = {"edge_index": data.edge_index.t().to(torch.long), "diff": unit, "dist": dist}
        if self.use_edge_attr:
            conv_args["edge_attr"] = data.edge_attr
        x = self._gps_embed(data, conv_args) if self.use_global_attn else data.x
        v = torch.zeros(x.size(0), 3, x.size(1))
        return x, v, conv_args


# -- models/PNAEqStack.py --------------------------------------------------------
def pnaeq_rbf(edge_dist, num_radial, cutoff):                                  # rbf_BasisLayer :453-493
    """sinc expansion times the cosine cutoff; edge_dist is [E]."""
    n = torch.arange(num_radial, device=edge_dist.device) + 1
    sinc = torch.sin(edge_dist.unsqueeze(-1) * n * torch.pi / cutoff) / edge_dist.unsqueeze(-1)
    cut = torch.where(edge_dist < cutoff, 0.5 * (torch.cos(torch.pi * edge_dist / cutoff) + 1),
                      torch.tensor(0.0, dtype=edge_dist.dtype))
    return sinc * cut.unsqueeze(-1)


class PNAEqMessage(nn.Module):                                                 # PainnMessage :224-395
    def __init__(self, node_size, x_aggregators, x_scalers, deg, edge_dim, num_radial):
        super().__init__()
        self.node_size, self.edge_dim = node_size, edge_dim
        F_in = node_size
        self.aggr_module = pyg.DegreeScalerAggregation(x_aggregators, x_scalers, deg)
        self.pre_nns = nn.ModuleList([nn.Sequential(nn.Linear((4 if edge_dim else 3) * F_in, F_in))])
        self.post_nns = nn.ModuleList([nn.Sequential(
            nn.Linear((len(x_aggregators) * len(x_scalers) + 1) * F_in, F_in))])
        self.rbf_emb = nn.Sequential(nn.Linear(num_radial, F_in), nn.Tanh())
        if edge_dim is not None:
            self.edge_encoder = nn.Linear(edge_dim, F_in)
        self.rbf_lin = nn.Linear(num_radial, F_in * 3, bias=False)
        self.scalar_message_mlp = nn.Sequential(nn.Linear(F_in, F_in), nn.Tanh(), nn.Linear(F_in, F_in),
                                                nn.SiLU(), nn.Linear(F_in, F_in * 3))

    def forward(self, x, v, edge_index, edge_rbf, edge_vec, edge_attr=None):   # :318-389
        src, dst = edge_index.t()
        parts = [x[src], x[dst], self.rbf_emb(edge_rbf)]
        if edge_attr is not None:
            parts.append(self.edge_encoder(edge_attr))
        scalar_out = self.scalar_message_mlp(self.pre_nns[0](torch.cat(parts, dim=-1)))
        filter_out = scalar_out * self.rbf_lin(edge_rbf)
        g_sv, g_ev, m_s = torch.split(filter_out, self.node_size, dim=-1)
        m_v = v[dst] * g_sv.unsqueeze(1) + g_ev.unsqueeze(1) * edge_vec.unsqueeze(-1)
        agg = self.aggr_module(m_s, src, x.shape[0])
        delta_x = self.post_nns[0](torch.cat([x, agg], dim=-1))
        delta_v = torch.zeros_like(v).index_add_(0, src, m_v)
        return x + delta_x, v + delta_v


class PNAEqUpdate(nn.Module):                                                  # PainnUpdate :398-450
    def __init__(self, node_size, last_layer=False):
        super().__init__()
        self.update_X = nn.Linear(node_size, node_size)
        self.update_V = nn.Linear(node_size, node_size)
        self.last_layer = last_layer
        self.update_mlp = nn.Sequential(nn.Linear(node_size * 2, node_size), nn.SiLU(),
                                        nn.Linear(node_size, node_size * (2 if last_layer else 3)))

    def forward(self, x, v):
        Xv, Vv = self.update_X(v), self.update_V(v)
        a = self.update_mlp(torch.cat((torch.linalg.norm(Vv, dim=1), x), dim=-1))
        inner = torch.sum(Xv * Vv, dim=1)
        if self.last_layer:
            a_xv, a_xx = torch.split(a, v.shape[-1], dim=-1)
            return x + a_xv * inner + a_xx
        a_vv, a_xv, a_xx = torch.split(a, x.shape[-1], dim=-1)
        return x + a_xv * inner + a_xx, v + a_vv.unsqueeze(1) * Xv


class PNAEqStack(Base):                                                        # :40-221
    def __init__(self, input_args, conv_args, deg, edge_dim, num_radial, radius, *args, **kwargs):
        self.x_aggregators = ["mean", "min", "max", "std"]
        self.x_scalers = ["identity", "amplification", "attenuation", "linear", "inverse_linear"]
        self.deg = torch.Tensor(deg)
        self.edge_dim, self.num_radial, self.radius = edge_dim, num_radial, radius
        self.is_edge_model = True
        super().__init__(input_args, conv_args, *args, **kwargs)

    def _init_conv(self):                                                      # :75-100
        dims = [self.embed_dim] + [self.hidden_dim] * self.num_conv_layers
        for k, (a, b) in enumerate(zip(dims[:-1], dims[1:])):
            last = k == self.num_conv_layers - 1
            self.graph_convs.append(self._apply_global_attn(self.get_conv(a, b, last, edge_dim=self.edge_embed_dim)))
            self.feature_layers.append(nn.Identity())

    def get_conv(self, input_dim, output_dim, last_layer=False, edge_dim=None):  # :102-176
        assert (output_dim if input_dim == 1 else input_dim) > 1
        msg = PNAEqMessage(input_dim, self.x_aggregators, self.x_scalers, self.deg, edge_dim, self.num_radial)
        upd = PNAEqUpdate(input_dim, last_layer)
        s_out = nn.Sequential(nn.Linear(input_dim, output_dim), nn.Tanh(), nn.Linear(output_dim, output_dim))
        both = "inv_node_feat, equiv_node_feat -> inv_node_feat, equiv_node_feat"
        steps = [(msg, self.conv_args + " -> inv_node_feat, equiv_node_feat")]
        if not last_layer:
            steps += [(upd, both), (s_out, "inv_node_feat -> inv_node_feat"),
                      (nn.Linear(input_dim, output_dim), "equiv_node_feat -> equiv_node_feat")]
        else:
            steps += [(upd, "inv_node_feat, equiv_node_feat -> inv_node_feat"),
                      (s_out, "inv_node_feat -> inv_node_feat")]
        steps.append((lambda inv_node_feat, equiv_node_feat: [inv_node_feat, equiv_node_feat], both))
        return pyg.Sequential(self.input_args, steps)

    def _embedding(self, data):                                                # :178-221
        Base._embedding(self, data)
        unit, dist = get_edge_vectors_and_lengths(data.pos, data.edge_index, data.edge_shifts, normalize=True)
        conv_args = {"edge_index": data.edge_index.t().to(torch.long),
                     "edge_rbf": pnaeq_rbf(dist.squeeze(), self.num_radial, self.radius), "edge_vec": unit}
        if self.use_edge_attr:
            conv_args["edge_attr"] = data.edge_attr
        x = self._gps_embed(data, conv_args) if self.use_global_attn else data.x
        v = torch.zeros(x.size(0), 3, x.size(1))
        return x, v, conv_args


# -- models/create.py:86-519 (the four stacks on the hot path) -------------------
def create_model(mpnn_type, input_dim, hidden_dim, output_dim, pe_dim, global_attn_engine,
                 global_attn_type, global_attn_heads, output_type, output_heads, activation_function,
                 loss_function_type, task_weights, num_conv_layers, freeze_conv=False, initial_bias=None,
                 num_nodes=None, max_neighbours=None, edge_dim=None, pna_deg=None, num_radial=None,
                 num_gaussians=None, num_filters=None, radius=None, equivariance=False,
                 envelope_exponent=None, **unused):
    torch.manual_seed(0)                                                       # :131
    common = (input_dim, hidden_dim, output_dim, pe_dim, global_attn_engine, global_attn_type,
              global_attn_heads, output_type, output_heads, activation_function, loss_function_type,
              equivariance)
    kw = dict(loss_weights=task_weights, freeze_conv=freeze_conv, initial_bias=initial_bias,
              num_conv_layers=num_conv_layers, num_nodes=num_nodes)
    if mpnn_type == "PNA":
        return PNAStack("inv_node_feat, equiv_node_feat, edge_index", "inv_node_feat, edge_index",
                        pna_deg, edge_dim, *common, **kw)
    if mpnn_type == "PNAEq":                                                    # create.py:446-470
        sig = "inv_node_feat, equiv_node_feat, edge_index, edge_rbf, edge_vec"
        return PNAEqStack(sig, sig, pna_deg, edge_dim, num_radial, radius, *common, **kw)
    if mpnn_type == "PNAPlus":                                                  # create.py:185-214
        return PNAPlusStack("inv_node_feat, equiv_node_feat, edge_index, rbf", "inv_node_feat, edge_index, rbf",
                            pna_deg, edge_dim, envelope_exponent, num_radial, radius, *common, **kw)
    if mpnn_type == "SchNet":
        return SCFStack("", "inv_node_feat, equiv_node_feat, edge_index, edge_weight, edge_rbf",
                        num_filters, edge_dim, num_gaussians, radius, *common,
                        max_neighbours=max_neighbours, **kw)
    if mpnn_type == "EGNN":
        return EGCLStack("inv_node_feat, equiv_node_feat, edge_index, edge_attr", "", edge_dim,
                         *common, max_neighbours=max_neighbours, **kw)
    if mpnn_type == "PAINN":
        sig = "inv_node_feat, equiv_node_feat, edge_index, diff, dist"
        return PAINNStack(sig, sig, edge_dim, num_radial, radius, *common, **kw)
    raise ValueError("Unknown mpnn_type: {0}".format(mpnn_type))


# -- preprocess/graph_samples_checks_and_updates.py:141-282 ----------------------
def radius_graph_pbc(pos, cell, pbc, r, max_num_neighbors, loop=False):
    """RadiusGraphPBC.__call__ without the retry / artificial-edge fallback (:178-221 are
    excluded from parity).  Returns (edge_index int64 [2,E], shift_int [E,3],
    length fp64 [E]) ordered by (dst, length) with ties broken by (src, S) -- the reference's
    np.lexsort((length, dst)) is stable over ASE's internal order, which is not specified."""
    i, j, d, S = pyg.neighbor_list_ase(pos, cell, pbc, r, self_interaction=True)
    if not loop:                                                               # :240-254
        keep = ~((i == j) & np.all(S == 0, axis=1))
        i, j, d, S = i[keep], j[keep], d[keep], S[keep]
    order = np.lexsort((S[:, 2], S[:, 1], S[:, 0], i, d, j))                   # :259
    i, j, d, S = i[order], j[order], d[order], S[order]
    if max_num_neighbors and max_num_neighbors > 0:                            # :263-274
        start = np.searchsorted(j, j, side="left")
        keep = (np.arange(j.size) - start) < max_num_neighbors
        i, j, d, S = i[keep], j[keep], d[keep], S[keep]
    return np.stack([i, j]).astype(np.int64), S.astype(np.int64), d


def radius_graph_pbc_full(pos, cell, pbc, r, max_num_neighbors, loop=False,
                          cutoff_multiplier=1.25, max_attempts=3):
    """RadiusGraphPBC.__call__ with its retry loop (:163-206) and the artificial-edge fallback
    `_ensure_connected` (:284-307).  The fallback draws from numpy's GLOBAL RNG exactly like the
    reference (np.random.choice per missing node, ascending node order), so a caller that seeds
    np.random the same way gets the same artificial sources.  Returns (edge_index, shift_int,
    length, cutoff_used)."""
    n = int(np.asarray(pos).shape[0])
    cutoff = float(r)
    for attempt in range(max_attempts):
        ei, S, d = radius_graph_pbc(pos, cell, pbc, cutoff, max_num_neighbors, loop)
        if np.unique(ei[1]).size == n:                                         # :196
            break
        if attempt < max_attempts - 1:                                         # :198-204
            cutoff *= cutoff_multiplier
            continue
        nodes = np.arange(n)                                                   # :284-307
        src, dst = list(ei[0]), list(ei[1])
        for m in np.setdiff1d(nodes, np.unique(ei[1])):
            s = np.random.choice(nodes[nodes != m]) if n > 1 else 0
            src.append(int(s)); dst.append(int(m))
        extra = len(src) - ei.shape[1]
        ei = np.stack([np.asarray(src), np.asarray(dst)]).astype(np.int64)
        S = np.concatenate([S, np.zeros((extra, 3), dtype=S.dtype)])
        d = np.concatenate([d, np.full(extra, cutoff - 1e-8)])
    return ei, S, d, cutoff
