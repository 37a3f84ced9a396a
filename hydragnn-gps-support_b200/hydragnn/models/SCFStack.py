"""SchNet stack on the B200 path.

Constructor, `get_conv`, `_embedding` and parameter names follow the reference's
SCFStack / CFConv (hydragnn/models/SCFStack.py:34-293) so configs and state dicts
drop in.  What runs underneath is different:

  * the aggregation  out[t] = sum_{e: dst(e)=t} lin1(x)[src(e)] * W_e  (PyG propagate,
    SCFStack.py:285,292-293) is ONE deterministic CSR kernel (ops.gather_mul_segment_sum)
    whose backward / double-backward are the sibling kernels of the contraction family;
  * the per-layer RadiusInteractionGraph of plain mode (SCFStack.py:59-61,120-153) is the
    cell-list kernel, built once per forward while `pos` is unchanged (the reference
    rebuilds the identical graph in every layer);
  * the coordinate update (SCFStack.py:244-252) is a segmented mean over the same CSR.
"""
import math
from typing import Optional

import torch
from torch import nn

from ..graph import topo_of
from ..nn import FactoredEdgeAttr, GaussianSmearing, LazyGaussians, Sequential, ShiftedSoftplus
from ..ops import gather_mul_segment_sum, radius_graph, segment_mean
from ..ops.filter import cfconv_fused, fused_filter_ok, schnet_filter
from ..ops.node import linear_fast
from ..utils.model import get_edge_vectors_and_lengths
from .Base import Base


def _passthrough(inv_node_feat, equiv_node_feat):
    return [inv_node_feat, equiv_node_feat]


_PASS = (_passthrough, "inv_node_feat, equiv_node_feat -> inv_node_feat, equiv_node_feat")


class RadiusInteractionGraph(nn.Module):
    """(pos, batch) -> (edge_index [2,E] grouped by centre, edge_weight [E] = |pos_i - pos_j|).

    torch_cluster semantics (strict d^2 < r^2 in fp32, same graph segment, no self loops,
    nearest `max_num_neighbors` per centre).  The graph is cached while the same `pos`
    tensor comes back (non-equivariant SchNet never moves the atoms)."""

    def __init__(self, cutoff=10.0, max_num_neighbors=32):
        super().__init__()
        self.cutoff, self.max_num_neighbors = cutoff, max_num_neighbors

    def forward(self, pos, batch):
        # cached on the tensor object itself (not on its address, which the allocator reuses)
        # ... and keyed on the tensor's version counter: an in-place update of the positions (an MD or
        # relaxation loop, noise augmentation) rebuilds the graph, as the reference does every call
        static = getattr(pos, "_hgk_rgraph_static", None)
        if static is not None:
            # a step captured as a CUDA graph (train/step_graph.py): the neighbour search ran before the
            # replay, its result sits in this buffer (padded to the bucket, grouped by centre)
            topo_of(static, pos.size(0), sorted_side=1)
            vec, length = get_edge_vectors_and_lengths(pos, static, None)
            return static, length.squeeze(-1)
        key = (self.cutoff, self.max_num_neighbors, pos._version)
        cached = getattr(pos, "_hgk_rgraph", None)
        if cached is None or cached[0] != key:
            ptr = getattr(batch, "_hgk_ptr", None)
            if ptr is None:
                counts = torch.bincount(batch)
                ptr = torch.zeros(counts.numel() + 1, dtype=torch.int32, device=pos.device)
                ptr[1:] = torch.cumsum(counts, 0)
            g = radius_graph(pos, ptr, self.cutoff, self.max_num_neighbors, loop=False)
            edge_index = g["edge_index"]
            topo = topo_of(edge_index, pos.size(0), sorted_side=1)
            topo.set_csr(1, g["rowptr"], None)
            cached = (key, edge_index)
            pos._hgk_rgraph = cached
        edge_index = cached[1]
        vec, length = get_edge_vectors_and_lengths(pos, edge_index, None)
        return edge_index, length.squeeze(-1)


class CFConv(nn.Module):
    """Continuous-filter convolution (SCFStack.py:214-293)."""

    def __init__(self, in_channels, out_channels, num_filters, nn_, cutoff, equivariant=False):
        super().__init__()
        self.lin1 = nn.Linear(in_channels, num_filters, bias=False)
        self.lin2 = nn.Linear(num_filters, out_channels)
        self.nn = nn_
        self.cutoff = cutoff
        self.equivariant = equivariant
        if equivariant:
            tail = nn.Linear(num_filters, 1, bias=False)
            nn.init.xavier_uniform_(tail.weight, gain=0.001)
            self.coord_mlp = nn.Sequential(nn.Linear(num_filters, num_filters), nn.ReLU(), tail)
        self.reset_parameters()

    def reset_parameters(self):
        nn.init.xavier_uniform_(self.lin1.weight)
        nn.init.xavier_uniform_(self.lin2.weight)
        self.lin2.bias.data.fill_(0)

    def _filter(self, edge_rbf, edge_attr):
        """Filter-generating network on cat[rbf, edge_attr] times the cosine envelope
        (SCFStack.py:268-272).  Fast path: one fused tensor-core kernel fed with the raw distances
        and the un-embedded edge attribute.  Otherwise torch, with a factored edge attribute still
        folded into the first Linear (K = num_gaussians + r instead of + hidden)."""
        first, last = self.nn[0], self.nn[2]
        if isinstance(edge_rbf, LazyGaussians):
            G = edge_rbf.offset.numel()
            factored = isinstance(edge_attr, FactoredEdgeAttr)
            raw = edge_attr.raw if factored else None
            if (edge_attr is None or factored) and fused_filter_ok(edge_rbf.dist, raw, first.out_features, G) \
                    and last.in_features == last.out_features == first.out_features:
                w0r = first.weight[:, G:] @ edge_attr.matrix if factored else None
                return schnet_filter(edge_rbf.dist, raw, edge_rbf.offset, edge_rbf.coeff, self.cutoff,
                                     first.weight[:, :G], w0r, first.bias, last.weight, last.bias)
            dist = edge_rbf.dist
            edge_rbf = edge_rbf.materialize()
        else:
            dist = None
        if isinstance(edge_attr, FactoredEdgeAttr):
            G = edge_rbf.size(-1)
            weight = torch.cat([first.weight[:, :G], first.weight[:, G:] @ edge_attr.matrix], dim=1)
            h = torch.nn.functional.linear(torch.cat([edge_rbf, edge_attr.raw], dim=-1), weight, first.bias)
            for layer in list(self.nn)[1:]:
                h = layer(h)
        else:
            feats = edge_rbf if edge_attr is None else torch.cat([edge_rbf, edge_attr], dim=-1)
            h = self.nn(feats)
        return h, dist

    def _fused_args(self, edge_rbf, edge_attr):
        """Arguments of the fused filter kernel when it applies (see `_filter`), else None."""
        first, last = self.nn[0], self.nn[2]
        if not isinstance(edge_rbf, LazyGaussians):
            return None
        G = edge_rbf.offset.numel()
        factored = isinstance(edge_attr, FactoredEdgeAttr)
        raw = edge_attr.raw if factored else None
        if not ((edge_attr is None or factored) and fused_filter_ok(edge_rbf.dist, raw, first.out_features, G)
                and last.in_features == last.out_features == first.out_features):
            return None
        w0r = first.weight[:, G:] @ edge_attr.matrix if factored else None
        return (edge_rbf.dist, raw, edge_rbf.offset, edge_rbf.coeff, self.cutoff, first.weight[:, :G], w0r,
                first.bias, last.weight, last.bias)

    def forward(self, x, pos, edge_index, edge_weight, edge_rbf, edge_attr=None):
        topo = topo_of(edge_index, x.size(0))
        if not self.equivariant:
            fused = self._fused_args(edge_rbf, edge_attr)
            if fused is not None:
                # filter network + envelope + aggregation as one autograd node: its backward forms the
                # filter gradient inside the kernels instead of writing and re-reading an [E, F] tensor
                x1 = linear_fast(x, self.lin1.weight)
                return linear_fast(cfconv_fused(x1, topo, *fused), self.lin2.weight, self.lin2.bias)
        W = self._filter(edge_rbf, edge_attr)
        if isinstance(W, tuple):     # unfused: apply the cosine envelope here (SCFStack.py:268)
            envelope = 0.5 * (torch.cos(edge_weight * math.pi / self.cutoff) + 1.0)
            W = W[0] * envelope.view(-1, 1)
        x = linear_fast(x, self.lin1.weight)         # tensor-core kernel at 128-multiple widths
        if self.equivariant:
            unit, _ = get_edge_vectors_and_lengths(pos, edge_index, None, normalize=True, eps=1.0)
            trans = torch.clamp(unit * self.coord_mlp(W), min=-100, max=100)
            pos = pos + segment_mean(trans, topo, 0)
        x = linear_fast(gather_mul_segment_sum(x, W, topo, side_out=1, side_in=0), self.lin2.weight, self.lin2.bias)
        if self.equivariant:
            return x, pos
        return x


class SCFStack(Base):
    def __init__(self, input_args, conv_args, num_filters: int, edge_dim: int, num_gaussians: int,
                 radius: float, *args, max_neighbours: Optional[int] = None, **kwargs):
        self.radius = radius
        self.max_neighbours = max_neighbours
        self.num_filters = num_filters
        self.edge_dim = edge_dim
        self.num_gaussians = num_gaussians
        self.is_edge_model = True
        super().__init__(input_args, conv_args, *args, **kwargs)

    def _edge_mode(self):
        return self.use_edge_attr or (self.use_global_attn and self.is_edge_model)

    def _init_conv(self):
        self.distance_expansion = GaussianSmearing(0.0, self.radius, self.num_gaussians)
        self.distance_expansion.lazy = True     # CFConv expands the distances itself (fused kernel)
        self.interaction_graph = RadiusInteractionGraph(self.radius, self.max_neighbours)
        widths = [self.embed_dim] + [self.hidden_dim] * self.num_conv_layers
        for k, (w_in, w_out) in enumerate(zip(widths[:-1], widths[1:])):
            conv = self.get_conv(w_in, w_out, k == self.num_conv_layers - 1, edge_dim=self.edge_embed_dim)
            self.graph_convs.append(self._apply_global_attn(conv))
            self.feature_layers.append(nn.Identity())

    def get_conv(self, input_dim, output_dim, last_layer=False, edge_dim=None):
        width = self.num_gaussians + (edge_dim or 0)
        filter_net = nn.Sequential(nn.Linear(width, self.num_filters), ShiftedSoftplus(),
                                   nn.Linear(self.num_filters, self.num_filters))
        moves_atoms = self.equivariance and not last_layer
        cfconv = CFConv(input_dim, output_dim, self.num_filters, filter_net, self.radius, moves_atoms)
        if self._edge_mode():
            return Sequential(
                "inv_node_feat, equiv_node_feat, edge_index, edge_weight, edge_rbf, batch, edge_attr",
                [(cfconv, self.conv_args + " -> inv_node_feat"), _PASS])
        steps = [(self.interaction_graph, "equiv_node_feat, batch -> edge_index, edge_weight"),
                 (self.distance_expansion, "edge_weight -> edge_rbf")]
        if moves_atoms:
            steps.append((cfconv, self.conv_args + " -> inv_node_feat, equiv_node_feat"))
        else:
            steps += [(cfconv, self.conv_args + " -> inv_node_feat"), _PASS]
        return Sequential("inv_node_feat, equiv_node_feat, batch", steps)

    def _embedding(self, data):
        super()._embedding(data)
        if self._edge_mode():
            if self.equivariance:
                raise Exception("For SchNet if using edge attributes or encodings, then "
                                "E(3)-equivariance cannot be ensured. Please disable equivariance "
                                "or edge attributes.")
            # the reference zeroes edge_shifts here (SCFStack.py:166-168): PBC images are dropped
            data.edge_shifts = None
            _, dist = get_edge_vectors_and_lengths(data.pos, data.edge_index, None)
            edge_weight = dist.squeeze(-1)
            conv_args = {"edge_index": data.edge_index, "edge_weight": edge_weight,
                         "edge_rbf": self.distance_expansion(edge_weight), "batch": data.batch}
            if self.use_edge_attr:
                conv_args["edge_attr"] = data.edge_attr
        else:
            conv_args = {"batch": data.batch}
            if getattr(data, "ptr", None) is not None:
                # graph offsets for the in-model radius graph without a device-side bincount (a sync)
                try:
                    data.batch._hgk_ptr = data.ptr.to(device=data.pos.device, dtype=torch.int32)
                except Exception:
                    pass
        if self.use_global_attn:
            return self._gps_embeddings(data, conv_args), data.pos, conv_args
        return data.x, data.pos, conv_args

    def __str__(self):
        return "SCFStack"
