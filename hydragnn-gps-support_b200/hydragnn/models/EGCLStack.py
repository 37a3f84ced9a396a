"""EGNN stack on the B200 path.

API and parameter names follow the reference's EGCLStack / E_GCL
(hydragnn/models/EGCLStack.py:22-298).  Underneath:

  * the first edge-MLP layer, which the reference applies to the materialised
    concatenation [x_row, x_col, |r|, e] of width 2C+1+De (EGCLStack.py:240-245), is
    evaluated in split-weight form: two N-row GEMMs (x W_row^T, x W_col^T), one fused
    gather-add kernel over the edges, plus the rank-1 |r| term and the edge-attr GEMM.
    The [E, 2C+1+De] tensor is never built;
  * `unsorted_segment_sum` / `unsorted_segment_mean` at edge_index[0]
    (EGCLStack.py:253,263-271,292-298) are deterministic CSR reductions.
"""
from typing import Optional

import torch
import torch.nn.functional as F
from torch import nn

from ..graph import topo_of
from ..nn import Sequential, edge_linear
from ..ops import edge_gather_add, segment_mean, segment_sum
from ..utils.model import get_edge_vectors_and_lengths
from .Base import Base


def _passthrough(inv_node_feat, equiv_node_feat):
    return [inv_node_feat, equiv_node_feat]


class E_GCL(nn.Module):
    """E(n)-equivariant graph convolution layer."""

    def __init__(self, input_channels, output_channels, hidden_channels, edge_attr_dim=0,
                 nodes_attr_dim=0, act_fn=None, recurrent=False, coords_weight=1.0, attention=False,
                 clamp=False, norm_diff=True, tanh=True, equivariant=False):
        super().__init__()
        act_fn = nn.ReLU() if act_fn is None else act_fn
        self.in_channels = input_channels
        self.edge_attr_dim = edge_attr_dim
        self.coords_weight = coords_weight
        self.equivariant = equivariant
        self.norm_diff, self.tanh, self.clamp = norm_diff, tanh, clamp
        if recurrent or attention or nodes_attr_dim:
            raise NotImplementedError("recurrent / attention / node attributes are never enabled by "
                                      "the reference's EGCLStack.get_conv (EGCLStack.py:71-80)")
        self.edge_mlp = nn.Sequential(
            nn.Linear(2 * input_channels + 1 + edge_attr_dim, hidden_channels), act_fn,
            nn.Linear(hidden_channels, hidden_channels), act_fn)
        self.node_mlp = nn.Sequential(
            nn.Linear(hidden_channels + input_channels, hidden_channels), act_fn,
            nn.Linear(hidden_channels, output_channels))
        if equivariant:
            tail = nn.Linear(hidden_channels, 1, bias=False)
            nn.init.xavier_uniform_(tail.weight, gain=0.001)
            layers = [nn.Linear(hidden_channels, hidden_channels), act_fn, tail]
            if tanh:
                layers.append(nn.Tanh())
            self.coord_mlp = nn.Sequential(*layers)

    def edge_model(self, x, radial, edge_attr, topo):
        lin0 = self.edge_mlp[0]
        C = self.in_channels
        w = lin0.weight
        p = F.linear(x, w[:, :C], lin0.bias)          # rows gathered at edge_index[0]
        q = F.linear(x, w[:, C:2 * C])                # rows gathered at edge_index[1]
        h = edge_gather_add(p, q, topo, 0, 1) + radial * w[:, 2 * C]
        if edge_attr is not None:
            h = h + edge_linear(edge_attr, w[:, 2 * C + 1:])
        for layer in list(self.edge_mlp)[1:]:
            h = layer(h)
        return h

    def forward(self, x, coord, edge_index, edge_attr, node_attr=None):
        topo = topo_of(edge_index, x.size(0))
        # the reference discards PBC shifts here (EGCLStack.py:275-277)
        unit, radial = get_edge_vectors_and_lengths(coord, edge_index, None, normalize=self.norm_diff, eps=1.0)
        m = self.edge_model(x, radial, edge_attr, topo)
        if self.equivariant:
            trans = torch.clamp(unit * self.coord_mlp(m), min=-100, max=100)
            coord = coord + segment_mean(trans, topo, 0) * self.coords_weight
        agg = segment_sum(m, topo, 0)
        lin0 = self.node_mlp[0]
        C = self.in_channels
        # node_mlp[0](cat[x, agg]) with the weight split, no [N, C+H] concat
        h = F.linear(x, lin0.weight[:, :C], lin0.bias) + F.linear(agg, lin0.weight[:, C:])
        for layer in list(self.node_mlp)[1:]:
            h = layer(h)
        if self.equivariant:
            return h, coord
        return h


class EGCLStack(Base):
    def __init__(self, input_args, conv_args, edge_attr_dim: int, *args,
                 max_neighbours: Optional[int] = None, **kwargs):
        self.edge_dim = 0 if edge_attr_dim is None else edge_attr_dim
        self.is_edge_model = True
        super().__init__(input_args, conv_args, *args, **kwargs)
        assert self.input_args == "inv_node_feat, equiv_node_feat, edge_index, edge_attr"

    def _init_conv(self):
        widths = [self.embed_dim] + [self.hidden_dim] * self.num_conv_layers
        for k, (w_in, w_out) in enumerate(zip(widths[:-1], widths[1:])):
            conv = self.get_conv(w_in, w_out, k == self.num_conv_layers - 1, edge_dim=self.edge_embed_dim)
            self.graph_convs.append(self._apply_global_attn(conv))
            self.feature_layers.append(nn.Identity())

    def get_conv(self, input_dim, output_dim, last_layer=False, edge_dim=None):
        if not edge_dim:
            edge_dim = self.edge_dim
        moves_atoms = self.equivariance and not last_layer
        egcl = E_GCL(input_channels=input_dim, output_channels=output_dim,
                     hidden_channels=self.hidden_dim, edge_attr_dim=edge_dim, equivariant=moves_atoms)
        sig = "inv_node_feat, equiv_node_feat, edge_index, edge_attr"
        if moves_atoms:
            return Sequential(self.input_args, [(egcl, sig + " -> inv_node_feat, equiv_node_feat")])
        return Sequential(self.input_args, [
            (egcl, sig + " -> inv_node_feat"),
            (_passthrough, "inv_node_feat, equiv_node_feat -> inv_node_feat, equiv_node_feat")])

    def _embedding(self, data):
        super()._embedding(data)
        data.edge_shifts = None   # EGCLStack.py:113-115 zeroes the shifts
        conv_args = {"edge_index": data.edge_index, "edge_attr": None}
        if self.edge_dim > 0:
            assert data.edge_attr is not None, "Data must have edge attributes if use_edge_attributes is set."
            conv_args["edge_attr"] = data.edge_attr
        if self.use_global_attn:
            return self._gps_embeddings(data, conv_args), data.pos, conv_args
        return data.x, data.pos, conv_args

    def __str__(self):
        return "EGCLStack"
