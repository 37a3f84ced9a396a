"""PaiNN stack on the B200 path.

Constructor / `get_conv` / `_embedding` contract and parameter names follow the reference's
PAINNStack, PainnMessage and PainnUpdate (hydragnn/models/PAINNStack.py:27-343).  The
message step, which the reference evaluates as index_select -> [E,3F] / [E,3,F] temporaries
-> two index_add_ (PAINNStack.py:237-257), is expressed as three calls of the CSR
contraction kernel -- every term is "gathered node row x per-edge filter (x per-edge
direction)", summed per receiving atom in a fixed order:

    ds[i]      = sum_e  phi_s [j_e,:]   * fw_s [e,:]
    dv[i,c,:]  = sum_e (v*phi_sv)[j_e,c,:] * fw_sv[e,:]  +  sum_e phi_ev[j_e,:] * fw_ev[e,:] * u[e,c]

with u = r_hat / d (the reference divides the unit vector by d once more, :247-251).
The kernels are closed under differentiation, which is what compute_grad_energy's
create_graph=True force loss needs (Base.py:614-620).
"""
import torch
from torch import nn

from ..graph import topo_of
from ..nn import Sequential, edge_linear
from ..ops import contract_node
from ..ops.node import linear_fast
from ..utils.model import get_edge_vectors_and_lengths
from .Base import Base


def _passthrough(inv_node_feat, equiv_node_feat):
    return [inv_node_feat, equiv_node_feat]


_BOTH = "inv_node_feat, equiv_node_feat -> inv_node_feat, equiv_node_feat"


def sinc_expansion(edge_dist: torch.Tensor, edge_size: int, cutoff: float):
    """sin(n pi d / cutoff) / d for n = 1..edge_size; edge_dist is [E,1]."""
    n = torch.arange(edge_size, device=edge_dist.device) + 1
    return torch.sin(edge_dist * n * torch.pi / cutoff) / edge_dist


def cosine_cutoff(edge_dist: torch.Tensor, cutoff: float):
    """0.5 (cos(pi d / cutoff) + 1) inside the cutoff, 0 outside."""
    inside = 0.5 * (torch.cos(torch.pi * edge_dist / cutoff) + 1)
    return torch.where(edge_dist < cutoff, inside, torch.zeros_like(inside))


def _mlp_fast(seq, x):
    """nn.Sequential of Linears and activations with the Linears on the tensor-core node-row kernel when the
    widths allow it (ops/node.py: linear_fast; torch otherwise)."""
    for layer in seq:
        x = linear_fast(x, layer.weight, layer.bias) if isinstance(layer, nn.Linear) else layer(x)
    return x


class PainnMessage(nn.Module):
    def __init__(self, node_size: int, edge_size: int, cutoff: float, edge_dim=None):
        super().__init__()
        self.node_size, self.edge_size, self.cutoff, self.edge_dim = node_size, edge_size, cutoff, edge_dim
        self.scalar_message_mlp = nn.Sequential(
            nn.Linear(node_size, node_size), nn.SiLU(), nn.Linear(node_size, node_size * 3))
        self.filter_layer = nn.Linear(edge_size, node_size * 3)
        if edge_dim is not None:
            self.edge_filter = nn.Sequential(
                nn.Linear(edge_dim, node_size), nn.SiLU(), nn.Linear(node_size, node_size * 3))

    def forward(self, node_scalar, node_vector, edge, edge_diff, edge_dist, edge_attr=None):
        Fs = self.node_size
        topo = getattr(edge, "_hgk_topo", None)
        if topo is None:   # `edge` is edge_index.T ([E,2]); receivers are column 0
            topo = topo_of(edge.t().contiguous(), node_scalar.size(0))
        # fw = (W sinc(d) + b) * fcut(d)  (PAINNStack.py:230-235) with the cutoff folded into the radial basis:
        # fw = [sinc(d) fcut(d) | fcut(d)] [W | b]^T.  The cutoff is then applied to R + 1 columns instead of
        # 3F, and neither the forward nor the (double) backward has an elementwise pass over [E, 3F].
        cut = cosine_cutoff(edge_dist, self.cutoff)
        basis = torch.cat((sinc_expansion(edge_dist, self.edge_size, self.cutoff) * cut, cut), dim=1)
        fw = torch.nn.functional.linear(
            basis, torch.cat((self.filter_layer.weight, self.filter_layer.bias.unsqueeze(1)), dim=1))
        if edge_attr is not None and self.edge_dim is not None:
            first = self.edge_filter[0]
            gate = edge_linear(edge_attr, first.weight, first.bias)
            for layer in list(self.edge_filter)[1:]:
                gate = layer(gate)
            fw = fw * gate
        phi = _mlp_fast(self.scalar_message_mlp, node_scalar)            # [N, 3F]
        fw_sv, fw_ev, fw_s = fw[:, :Fs], fw[:, Fs:2 * Fs], fw[:, 2 * Fs:]
        phi_sv, phi_ev, phi_s = phi[:, :Fs], phi[:, Fs:2 * Fs], phi[:, 2 * Fs:]
        u = edge_diff / edge_dist                                        # [E,3]
        ds = contract_node(topo, 0, A=phi_s, side_a=1, W=fw_s, C=1, F=Fs)
        dv = contract_node(topo, 0, A=phi_ev, side_a=1, W=fw_ev, U=u, C=3, CO=3, F=Fs)
        if node_vector.requires_grad or bool(getattr(node_vector, "_hgk_nonzero", True)):
            dv = dv + contract_node(topo, 0, A=node_vector * phi_sv.unsqueeze(1), side_a=1,
                                    W=fw_sv, C=3, CO=3, F=Fs)
        return node_scalar + ds, node_vector + dv


class PainnUpdate(nn.Module):
    """Node-level update (GEMM-bound, stays in torch): U/V on the vectors, gated mixing."""

    def __init__(self, node_size: int, last_layer=False):
        super().__init__()
        self.update_U = nn.Linear(node_size, node_size)
        self.update_V = nn.Linear(node_size, node_size)
        self.last_layer = last_layer
        width_out = node_size * (2 if last_layer else 3)
        self.update_mlp = nn.Sequential(
            nn.Linear(node_size * 2, node_size), nn.SiLU(), nn.Linear(node_size, width_out))

    def forward(self, node_scalar, node_vector):
        flat = node_vector.reshape(-1, node_vector.shape[-1])            # [3N, F] rows for the tensor-core Linear
        Uv = linear_fast(flat, self.update_U.weight, self.update_U.bias).view_as(node_vector)
        Vv = linear_fast(flat, self.update_V.weight, self.update_V.bias).view_as(node_vector)
        gates = _mlp_fast(self.update_mlp, torch.cat((torch.linalg.norm(Vv, dim=1), node_scalar), dim=1))
        inner = torch.sum(Uv * Vv, dim=1)
        width = node_vector.shape[-1]
        if self.last_layer:
            a_sv, a_ss = torch.split(gates, width, dim=1)
            return node_scalar + a_sv * inner + a_ss
        a_vv, a_sv, a_ss = torch.split(gates, width, dim=1)
        return node_scalar + a_sv * inner + a_ss, node_vector + a_vv.unsqueeze(1) * Uv


class PAINNStack(Base):
    def __init__(self, input_args, conv_args, edge_dim: int, num_radial: int, radius: float,
                 *args, **kwargs):
        self.edge_dim = edge_dim
        self.num_radial = num_radial
        self.radius = radius
        self.is_edge_model = True
        super().__init__(input_args, conv_args, *args, **kwargs)

    def _init_conv(self):
        widths = [self.embed_dim] + [self.hidden_dim] * self.num_conv_layers
        for k, (w_in, w_out) in enumerate(zip(widths[:-1], widths[1:])):
            conv = self.get_conv(w_in, w_out, k == self.num_conv_layers - 1, edge_dim=self.edge_embed_dim)
            self.graph_convs.append(self._apply_global_attn(conv))
            self.feature_layers.append(nn.Identity())

    def get_conv(self, input_dim, output_dim, last_layer=False, edge_dim=None):
        width = output_dim if input_dim == 1 else input_dim
        assert width > 1, "PainnNet requires more than one hidden dimension between input_dim and output_dim."
        message = PainnMessage(node_size=input_dim, edge_size=self.num_radial, cutoff=self.radius,
                               edge_dim=edge_dim)
        update = PainnUpdate(node_size=input_dim, last_layer=last_layer)
        scalar_out = nn.Sequential(nn.Linear(input_dim, output_dim), nn.Tanh(),
                                   nn.Linear(output_dim, output_dim))
        steps = [(message, self.conv_args + " -> inv_node_feat, equiv_node_feat")]
        if last_layer:
            steps += [(update, "inv_node_feat, equiv_node_feat -> inv_node_feat"),
                      (scalar_out, "inv_node_feat -> inv_node_feat")]
        else:
            steps += [(update, _BOTH), (scalar_out, "inv_node_feat -> inv_node_feat"),
                      (nn.Linear(input_dim, output_dim), "equiv_node_feat -> equiv_node_feat")]
        steps.append((_passthrough, _BOTH))
        return Sequential(self.input_args, steps)

    def _embedding(self, data):
        super()._embedding(data)
        assert data.pos is not None, "PAINN requires node positions (data.pos) to be set."
        unit, dist = get_edge_vectors_and_lengths(data.pos, data.edge_index, data.edge_shifts, normalize=True)
        edge = data.edge_index.t()
        edge._hgk_topo = topo_of(data.edge_index, data.pos.size(0))
        conv_args = {"edge_index": edge, "diff": unit, "dist": dist}
        if self.use_edge_attr:
            assert data.edge_attr is not None, "Data must have edge attributes if use_edge_attributes is set."
            conv_args["edge_attr"] = data.edge_attr
        x = self._gps_embeddings(data, conv_args) if self.use_global_attn else data.x
        v = torch.zeros(x.size(0), 3, x.size(1), device=x.device)
        v._hgk_nonzero = False   # first layer: the v*phi term of the message is identically zero
        return x, v, conv_args

    def __str__(self):
        return "PAINNStack"
