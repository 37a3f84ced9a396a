"""Common encoder / decoder of every stack -- B200 path.

Keeps the constructor, attributes and method contract of the reference's Base
(hydragnn/models/Base.py:31-686): `_init_conv`, `get_conv`, `_embedding(data) ->
(x, pos|v, conv_args)`, `forward(data) -> list of head outputs`, `loss`,
`loss_hpweighted`, `energy_force_loss`; attributes graph_convs, feature_layers,
heads_NN, graph_shared, num_heads, head_dims, head_type, loss_weights, var_output
are the ones the train loop reads (train/train_validate_test.py:94-96,320-359).

What changed underneath: pooling (global_mean_pool, :478) and the per-graph energy
sum (torch_scatter.scatter_add, :598) are deterministic segmented reductions of
the contraction family, fed by the batch's `ptr` (no `.unique()` device syncs,
:484-485), and every edge gather/scatter inside the convs goes through a Topo
(CSR) cached on the batch.
"""
import inspect

import torch
from torch.nn import Linear, Module, ModuleDict, ModuleList, Sequential

from ..globalAtt.gps import GPSConv
from ..graph import pool_topo
from ..nn import BatchNorm, FactoredEdgeAttr
from ..ops import segment_sum
from ..ops.node import linear_fast
from ..utils.model import activation_function_selection, loss_function_selection


_WHOLE_BATCH = {}


def _whole_batch_segment(num_nodes, device):
    """cu_seqlens = [0, N] on `device`, cached: torch.tensor(..., device=cuda) is a blocking copy
    from pageable memory that drains the stream -- once per forward it kept the host from running
    ahead of the GPU."""
    key = (int(num_nodes), str(device))
    cu = _WHOLE_BATCH.get(key)
    if cu is None:
        if len(_WHOLE_BATCH) > 64:
            _WHOLE_BATCH.clear()
        cu = torch.tensor([0, int(num_nodes)], dtype=torch.int32, device=device)
        _WHOLE_BATCH[key] = cu
    return cu.clone()      # device-side copy: callers attach attributes to it


class Base(Module):
    def __init__(
        self,
        input_args: str,
        conv_args: str,
        input_dim: int,
        hidden_dim: int,
        output_dim: list,
        pe_dim: int,
        global_attn_engine: str,
        global_attn_type: str,
        global_attn_heads: int,
        output_type: list,
        config_heads: dict,
        activation_function_type: str,
        loss_function_type: str,
        equivariance: bool,
        ilossweights_hyperp: int = 1,
        loss_weights: list = [1.0, 1.0, 1.0],
        ilossweights_nll: int = 0,
        freeze_conv=False,
        initial_bias=None,
        dropout: float = 0.25,
        num_conv_layers: int = 16,
        num_nodes: int = None,
    ):
        super().__init__()
        self.input_args, self.conv_args = input_args, conv_args
        self.global_attn_engine, self.global_attn_type = global_attn_engine, global_attn_type
        self.input_dim, self.pe_dim, self.hidden_dim = input_dim, pe_dim, hidden_dim
        self.global_attn_heads = global_attn_heads
        self.dropout = self.global_attn_dropout = dropout
        self.num_conv_layers, self.num_nodes = num_conv_layers, num_nodes
        self.graph_convs, self.feature_layers, self.heads_NN = ModuleList(), ModuleList(), ModuleList()
        self.config_heads, self.head_type, self.head_dims = config_heads, output_type, output_dim
        self.num_heads = len(self.head_dims)
        self.convs_node_hidden, self.batch_norms_node_hidden = ModuleDict({}), ModuleDict({})
        self.convs_node_output, self.batch_norms_node_output = ModuleDict({}), ModuleDict({})
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
                raise ValueError("Inconsistent number of loss weights and tasks: "
                                 + str(len(loss_weights)) + " VS " + str(self.num_heads))
            norm = sum(abs(w) for w in loss_weights)
            self.loss_weights = [w / norm for w in loss_weights]

        self.use_edge_attr = False
        if getattr(self, "edge_dim", None) is not None and self.edge_dim > 0:
            self.use_edge_attr = True
            self._require_edge_attr_arg()

        self.freeze_conv, self.initial_bias = freeze_conv, initial_bias

        if self.global_attn_engine:
            self.use_global_attn = True
            self.embed_dim = self.edge_embed_dim = hidden_dim
            if self.is_edge_model:
                self._require_edge_attr_arg()
        else:
            self.use_global_attn = False
            self.embed_dim = input_dim
            self.edge_embed_dim = getattr(self, "edge_dim", None)
        self.use_encodings = False

        if self.use_global_attn or self.use_encodings:
            self.pos_emb = Linear(self.pe_dim, self.hidden_dim, bias=False)
            if self.input_dim:
                self.node_emb = Linear(self.input_dim, self.hidden_dim, bias=False)
                self.node_lin = Linear(2 * self.hidden_dim, self.hidden_dim, bias=False)
            if self.is_edge_model:
                self.rel_pos_emb = Linear(self.pe_dim, self.hidden_dim, bias=False)
                if self.use_edge_attr:
                    self.edge_emb = Linear(self.edge_dim, self.hidden_dim, bias=False)
                    self.edge_lin = Linear(2 * self.hidden_dim, self.hidden_dim, bias=False)

        self._init_conv()
        if self.freeze_conv:
            self._freeze_conv()
        self._multihead()
        if self.initial_bias is not None:
            self._set_bias()
        self.conv_checkpointing = False

    # ------------------------------------------------------------- wiring ---
    def _require_edge_attr_arg(self):
        if "edge_attr" not in self.input_args:
            self.input_args += ", edge_attr"
        if "edge_attr" not in self.conv_args:
            self.conv_args += ", edge_attr"

    def _apply_global_attn(self, mpnn):
        if self.use_global_attn:
            if self.global_attn_engine == "GPS":
                return GPSConv(channels=self.hidden_dim, conv=mpnn, heads=self.global_attn_heads,
                               dropout=self.global_attn_dropout, attn_type=self.global_attn_type)
            return None
        return mpnn

    def _init_conv(self):
        widths = [self.embed_dim] + [self.hidden_dim] * self.num_conv_layers
        for w_in, w_out in zip(widths[:-1], widths[1:]):
            self.graph_convs.append(
                self._apply_global_attn(self.get_conv(w_in, w_out, edge_dim=self.edge_embed_dim)))
            self.feature_layers.append(BatchNorm(self.hidden_dim))

    def _gps_embeddings(self, data, conv_args):
        """Node / edge input embeddings when global attention is on (Base.py:229-243)."""
        x = self.pos_emb(data.pe)
        if self.input_dim:
            # [N, 2 hidden] -> hidden on the tensor cores when the widths allow it (ops/node.py: linear_fast)
            x = linear_fast(torch.cat((self.node_emb(data.x.float()), x), 1), self.node_lin.weight, self.node_lin.bias)
        if self.is_edge_model:
            # e = rel_pos_emb(rel_pe)  [or edge_lin(cat(edge_emb(edge_attr), rel_pos_emb(rel_pe)))],
            # all bias-free Linears: handed to the convs in factored form (see nn.FactoredEdgeAttr)
            if self.use_edge_attr:
                c = self.hidden_dim
                w = self.edge_lin.weight
                raw = torch.cat((conv_args["edge_attr"].float(), data.rel_pe), 1)
                matrix = torch.cat((w[:, :c] @ self.edge_emb.weight, w[:, c:] @ self.rel_pos_emb.weight), 1)
            else:
                raw, matrix = data.rel_pe, self.rel_pos_emb.weight
            conv_args["edge_attr"] = FactoredEdgeAttr(raw, matrix)
        return x

    def _embedding(self, data):
        if getattr(data, "edge_shifts", None) is None:
            data.edge_shifts = torch.zeros((data.edge_index.size(1), 3), device=data.edge_index.device)
        conv_args = {"edge_index": data.edge_index}
        if self.use_edge_attr:
            assert data.edge_attr is not None, \
                "Data must have edge attributes if use_edge_attributes is set."
            conv_args["edge_attr"] = data.edge_attr
        if self.use_global_attn:
            return self._gps_embeddings(data, conv_args), data.pos, conv_args
        return data.x, data.pos, conv_args

    def _freeze_conv(self):
        for group in (self.graph_convs, self.feature_layers):
            for param in group.parameters():
                param.requires_grad = False

    def _set_bias(self):
        for head, kind in zip(self.heads_NN, self.head_type):
            if kind == "graph":
                for branch in head.values():
                    branch[-1].bias.data.fill_(self.initial_bias)

    def _make_node_conv(self, w_in, w_out, last_layer):
        if "last_layer" in inspect.signature(self.get_conv).parameters:
            return self.get_conv(w_in, w_out, last_layer=last_layer)
        return self.get_conv(w_in, w_out)

    def _init_node_conv(self):
        branches = self.config_heads["node"]
        assert self.num_branches == len(branches) or self.num_branches == 1, \
            "asumming node head has the same branches as graph head, if any"
        if any(b["architecture"]["type"] != "conv" for b in branches):
            return
        node_heads = [i for i, kind in enumerate(self.head_type) if kind == "node"]
        if not node_heads:
            return
        for b in branches:
            arch, name = b["architecture"], b["type"]
            dims = arch["dim_headlayers"]
            hidden, hidden_bn, outs, outs_bn = ModuleList(), ModuleList(), ModuleList(), ModuleList()
            hidden.append(self._make_node_conv(self.hidden_dim, dims[0], False))
            hidden_bn.append(BatchNorm(dims[0]))
            for k in range(arch["num_headlayers"] - 1):
                hidden.append(self._make_node_conv(dims[k], dims[k + 1], False))
                hidden_bn.append(BatchNorm(dims[k + 1]))
            for ihead in node_heads:
                width = self.head_dims[ihead] * (1 + self.var_output)
                outs.append(self._make_node_conv(dims[-1], width, True))
                outs_bn.append(BatchNorm(width))
            self.convs_node_hidden[name], self.batch_norms_node_hidden[name] = hidden, hidden_bn
            self.convs_node_output[name], self.batch_norms_node_output[name] = outs, outs_bn

    def _multihead(self):
        self.graph_shared = ModuleDict({})
        self.num_branches = 1
        if "graph" in self.config_heads:
            self.num_branches = len(self.config_heads["graph"])
            for b in self.config_heads["graph"]:
                arch = b["architecture"]
                width = arch["dim_sharedlayers"]
                layers = [Linear(self.hidden_dim, width), self.activation_function]
                for _ in range(arch["num_sharedlayers"] - 1):
                    layers += [Linear(width, width), self.activation_function]
                self.graph_shared[b["type"]] = Sequential(*layers)
        if "node" in self.config_heads:
            self._init_node_conv()

        inode_feature = 0
        for ihead in range(self.num_heads):
            head_NN = ModuleDict({})
            width_out = self.head_dims[ihead] * (1 + self.var_output)
            if self.head_type[ihead] == "graph":
                for b in self.config_heads["graph"]:
                    arch = b["architecture"]
                    dh = arch["dim_headlayers"]
                    layers = [Linear(arch["dim_sharedlayers"], dh[0]), self.activation_function]
                    for k in range(arch["num_headlayers"] - 1):
                        layers += [Linear(dh[k], dh[k + 1]), self.activation_function]
                    layers.append(Linear(dh[-1], width_out))
                    head_NN[b["type"]] = Sequential(*layers)
            elif self.head_type[ihead] == "node":
                for b in self.config_heads["node"]:
                    arch, name = b["architecture"], b["type"]
                    kind = arch["type"]
                    if kind in ("mlp", "mlp_per_node"):
                        self.num_mlp = 1 if kind == "mlp" else self.num_nodes
                        assert self.num_nodes is not None, "num_nodes must be positive integer for MLP"
                        head_NN[name] = MLPNode(self.hidden_dim, width_out, self.num_mlp,
                                                arch["dim_headlayers"], kind, self.activation_function)
                    elif kind == "conv":
                        chain = ModuleList()
                        for conv, bn in zip(self.convs_node_hidden[name], self.batch_norms_node_hidden[name]):
                            chain.append(conv)
                            chain.append(bn)
                        chain.append(self.convs_node_output[name][inode_feature])
                        chain.append(self.batch_norms_node_output[name][inode_feature])
                        head_NN[name] = chain
                        inode_feature += 1
                    else:
                        raise ValueError(
                            "Unknown head NN structure for node features" + kind
                            + "; currently only support 'mlp', 'mlp_per_node' or 'conv'")
            else:
                raise ValueError("Unknown head type" + self.head_type[ihead]
                                 + "; currently only support 'graph' or 'node'")
            self.heads_NN.append(head_NN)

    def enable_conv_checkpointing(self):
        self.conv_checkpointing = True

    # ------------------------------------------------------------ forward ---
    def _pool_mean(self, x, data):
        topo = pool_topo(data, x.size(0))
        total = segment_sum(x, topo, 0)
        count = topo.degree(0).clamp(min=1).to(x.dtype).unsqueeze(-1)
        return total / count

    def _run_node_conv_head(self, chain, x, equiv, conv_args):
        conv_args = {k: v for k, v in conv_args.items() if k != "cu_seqlens"}
        for conv, bn in zip(chain[0::2], chain[1::2]):
            x, equiv = conv(inv_node_feat=x, equiv_node_feat=equiv, **conv_args)
            x = self.activation_function(bn(x))
        return x

    # "batch": all nodes of the mini-batch attend to each other, which is what the reference
    # computes (its GPSConv never receives graph_batch); "graph": per-graph segments.
    gps_segments = "batch"

    def _attention_segments(self, data, num_nodes, device):
        """cu_seqlens for the GPS layers, built once per forward (with the host-side sum of
        squared segment lengths attached for the kernel's flop accounting -- no device sync)."""
        static = getattr(data, "gps_cu_seqlens", None)
        if static is not None:          # a batch padded to its size bucket (train/step_graph.py): [0, N, N_b]
            return static
        bounds = None
        if self.gps_segments == "graph" and getattr(data, "ptr", None) is not None:
            cu = data.ptr.to(device=device, dtype=torch.int32)
            sum_sq = getattr(data, "sum_sq_nodes", None)
            sizes = getattr(data, "graph_sizes", None)              # host copy kept by Batch.from_data_list
            if sizes is not None:
                bounds = [0]
                for n in sizes:
                    bounds.append(bounds[-1] + int(n))
                sum_sq = sum(int(n) * int(n) for n in sizes)
            elif not data.ptr.is_cuda:
                bounds = [int(v) for v in data.ptr.tolist()]
        else:
            cu = _whole_batch_segment(num_nodes, device)
            sum_sq = num_nodes * num_nodes
            bounds = [0, int(num_nodes)]
        if sum_sq is not None:
            cu._hgk_sum_sq = int(sum_sq)
        if bounds is not None:
            cu._hgk_bounds = bounds        # host-side segment boundaries: no sync in the differentiable backward
        return cu

    def forward(self, data):
        pos = getattr(data, "pos", None)
        if self.use_global_attn and pos is not None and pos.requires_grad:
            # force loss: backward itself is differentiated, which a captured backward cannot provide
            from ..globalAtt import graphed_tail
            with graphed_tail.eager():
                return self._forward(data)
        return self._forward(data)

    def _forward(self, data):
        inv_node_feat, equiv_node_feat, conv_args = self._embedding(data)
        if self.use_global_attn:
            conv_args["cu_seqlens"] = self._attention_segments(
                data, inv_node_feat.size(0), inv_node_feat.device)
        for conv, feat_layer in zip(self.graph_convs, self.feature_layers):
            if self.conv_checkpointing:
                from torch.utils.checkpoint import checkpoint
                inv_node_feat, equiv_node_feat = checkpoint(
                    conv, use_reentrant=False, inv_node_feat=inv_node_feat,
                    equiv_node_feat=equiv_node_feat, **conv_args)
            else:
                inv_node_feat, equiv_node_feat = conv(
                    inv_node_feat=inv_node_feat, equiv_node_feat=equiv_node_feat, **conv_args)
            inv_node_feat = self.activation_function(feat_layer(inv_node_feat))
        x = inv_node_feat

        if getattr(data, "batch", None) is None:
            data.batch = torch.zeros(x.size(0), dtype=torch.long, device=x.device)
        x_graph = self._pool_mean(x, data)

        multi = self.num_branches > 1
        if multi:
            ids = data.dataset_name.reshape(-1)
            branch_ids = sorted(set(ids.tolist()))
            counts = pool_topo(data, x.size(0)).degree(0).long()
        outputs, outputs_var = [], []
        for head_dim, headloc, type_head in zip(self.head_dims, self.heads_NN, self.head_type):
            if type_head == "graph":
                if not multi:
                    out = headloc["branch-0"](self.graph_shared["branch-0"](x_graph))
                else:
                    out = x.new_zeros((x_graph.size(0), head_dim * (1 + self.var_output)))
                    for bid in branch_ids:
                        mask = ids == bid
                        name = f"branch-{bid}"
                        out[mask] = headloc[name](self.graph_shared[name](x_graph[mask]))
            else:
                node_kind = self.config_heads["node"][0]["architecture"]["type"]
                if not multi:
                    if node_kind == "conv":
                        out = self._run_node_conv_head(headloc["branch-0"], x, equiv_node_feat, conv_args)
                    else:
                        out = headloc["branch-0"](x=x, batch=data.batch)
                else:
                    out = x.new_zeros((x.size(0), head_dim * (1 + self.var_output)))
                    for bid in branch_ids:
                        mask_nodes = torch.repeat_interleave(ids == bid, counts)
                        name = f"branch-{bid}"
                        if node_kind == "conv":
                            raise NotImplementedError("multi-branch conv node heads are not on the B200 path")
                        out[mask_nodes] = headloc[name](x=x[mask_nodes], batch=data.batch[mask_nodes])
            outputs.append(out[:, :head_dim])
            outputs_var.append(out[:, head_dim:] ** 2)
        if self.var_output:
            return outputs, outputs_var
        return outputs

    # ------------------------------------------------------------- losses ---
    def loss(self, pred, value, head_index):
        var = None
        if self.var_output:
            var = pred[1]
            pred = pred[0]
        if self.ilossweights_nll == 1:
            raise ValueError("loss_nll() not ready yet")
        return self.loss_hpweighted(pred, value, head_index, var=var)

    def energy_force_loss(self, pred, data):
        assert data.pos is not None and data.energy is not None and data.forces is not None, \
            "data.pos, data.energy, data.forces must be provided for energy-force loss."
        assert data.pos.requires_grad, \
            "data.pos does not have grad, so force predictions cannot be computed."
        assert self.num_heads == 1 and self.head_type[0] == "node", \
            "Force predictions are only supported for models with one head that predict nodal energy."
        node_energy_pred = pred[0]
        topo = pool_topo(data, node_energy_pred.size(0))
        graph_energy_pred = segment_sum(node_energy_pred, topo, 0).squeeze().float()
        graph_energy_true = data.energy.squeeze().float()
        energy_loss_weight = self.loss_weights[0]
        energy_loss = self.loss_function(graph_energy_pred, graph_energy_true)
        tot_loss = energy_loss * energy_loss_weight
        forces_true = data.forces.float()
        forces_pred = torch.autograd.grad(
            graph_energy_pred, data.pos, grad_outputs=torch.ones_like(graph_energy_pred),
            retain_graph=graph_energy_pred.requires_grad, create_graph=True)[0].float()
        assert forces_pred is not None, "No gradients were found for data.pos."
        forces_pred = -forces_pred
        force_loss_weight = (energy_loss_weight * torch.mean(torch.abs(graph_energy_true))
                             / (torch.mean(torch.abs(forces_true)) + 1e-8))
        tot_loss = tot_loss + self.loss_function(forces_pred, forces_true) * force_loss_weight
        self.last_forces_pred = forces_pred
        return tot_loss, [energy_loss]

    def loss_hpweighted(self, pred, value, head_index, var=None):
        tot_loss = 0
        tasks_loss = []
        for ihead in range(self.num_heads):
            head_pre = pred[ihead]
            head_val = value[head_index[ihead]]
            if head_pre.shape != head_val.shape:
                head_val = torch.reshape(head_val, head_pre.shape)
            if var is None:
                assert self.loss_function_type != "GaussianNLLLoss", \
                    "Expecting var for GaussianNLLLoss, but got None"
                term = self.loss_function(head_pre, head_val)
            else:
                term = self.loss_function(head_pre, head_val, var[ihead])
            tot_loss = tot_loss + term * self.loss_weights[ihead]
            tasks_loss.append(term)
        return tot_loss, tasks_loss

    def __str__(self):
        return "Base"


class MLPNode(Module):
    """Node-level MLP head: one shared MLP ("mlp") or one per node slot ("mlp_per_node")."""

    def __init__(self, input_dim, output_dim, num_mlp, hidden_dim_node, node_type, activation_function):
        super().__init__()
        self.input_dim, self.output_dim, self.node_type = input_dim, output_dim, node_type
        self.num_mlp = self.num_nodes = num_mlp
        self.activation_function = activation_function
        self.mlp = ModuleList()
        for _ in range(self.num_mlp):
            widths = [input_dim] + list(hidden_dim_node)
            layers = []
            for a, b in zip(widths[:-1], widths[1:]):
                layers += [Linear(a, b), activation_function]
            layers.append(Linear(widths[-1], output_dim))
            self.mlp.append(Sequential(*layers))

    def forward(self, x: torch.Tensor, batch: torch.Tensor):
        if self.node_type == "mlp":
            return self.mlp[0](x)
        outs = x.new_zeros((x.shape[0], self.output_dim))
        for inode in range(self.num_nodes):
            rows = torch.arange(inode, batch.shape[0], self.num_nodes, device=x.device)
            outs[rows] = self.mlp[inode](x[rows])
        return outs

    def __str__(self):
        return "MLPNode"
