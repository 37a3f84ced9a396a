"""PNAEq stack on the B200 path (PaiNN-style equivariant messages with PNA aggregation of the scalar
channel) -- the model of the reference's physics-informed multibranch configuration
(examples/multibranch/gfm_multibranch_physics-informed.json:7,74).

Constructor / `get_conv` / `_embedding` contract and parameter names follow the reference's
PNAEqStack, PainnMessage, PainnUpdate and rbf_BasisLayer (hydragnn/models/PNAEqStack.py:40-493).
Per edge e = (i <- j) with i = edge_index[0] (where the reference aggregates) and j = edge_index[1]:

    m0_e   = pre([x_i, x_j, rbf_emb(rbf_e)(, edge_encoder(edge_attr_e))])     split-weight form: two
             N-row GEMMs + hgk_edge_gather_add, the E-row terms added on top
    f_e    = scalar_message_mlp(m0_e) * rbf_lin(rbf_e)  -> (g_sv | g_ev | m_s)    E-row GEMMs (torch)
    x_i   += post([x_i, scale(PNA-aggregate_i(m_s))])                          hgk_pna_aggregate (R only)
    v_i   += sum_e v_j * g_sv + g_ev (x) r_hat                                 two hgk_contract_node calls

All ops are closed under differentiation (the PNA aggregate switches to differentiable torch ops
when its backward is itself differentiated), which the force loss of that configuration needs.
"""
import torch
import torch.nn.functional as F
from torch import nn

from ..graph import topo_of
from ..nn import Sequential, edge_linear
from ..ops import contract_node, edge_gather_add, pna_aggregate
from ..utils.model import get_edge_vectors_and_lengths
from .Base import Base
from .PNAStack import PNAConv as _PNAScalers

_BOTH = "inv_node_feat, equiv_node_feat -> inv_node_feat, equiv_node_feat"


def _passthrough(inv_node_feat, equiv_node_feat):
    return [inv_node_feat, equiv_node_feat]


class rbf_BasisLayer(nn.Module):
    """sin(n pi d / cutoff) / d times the cosine cutoff (PNAEqStack.py:453-493); no parameters."""

    def __init__(self, num_radial: int, cutoff: float):
        super().__init__()
        self.num_radial, self.cutoff = num_radial, cutoff

    def forward(self, edge_dist):
        n = torch.arange(self.num_radial, device=edge_dist.device) + 1
        d = edge_dist.unsqueeze(-1)
        sinc = torch.sin(d * n * torch.pi / self.cutoff) / d
        cut = torch.where(edge_dist < self.cutoff, 0.5 * (torch.cos(torch.pi * edge_dist / self.cutoff) + 1),
                          torch.zeros_like(edge_dist))
        return sinc * cut.unsqueeze(-1)


class PainnMessage(nn.Module):
    def __init__(self, node_size, x_aggregators, x_scalers, deg, edge_dim, num_radial, towers=1, pre_layers=1,
                 post_layers=1, divide_input=False, **kwargs):
        super().__init__()
        if towers != 1 or pre_layers != 1 or post_layers != 1 or divide_input:
            raise NotImplementedError("only the reference's PNAEq configuration is on the B200 path")
        if list(x_aggregators) != ["mean", "min", "max", "std"]:
            raise NotImplementedError(f"aggregators {x_aggregators}: the kernel computes mean,min,max,std")
        self.node_size, self.num_radial, self.edge_dim = node_size, num_radial, edge_dim
        Fin = node_size
        self.scalers = list(x_scalers)
        hist = torch.as_tensor(deg, dtype=torch.float)
        bins = torch.arange(hist.numel(), dtype=torch.float)
        total = int(hist.sum())
        self.avg_deg_lin = float((bins * hist).sum()) / total
        self.avg_deg_log = float(((bins + 1).log() * hist).sum()) / total
        self.pre_nns = nn.ModuleList([nn.Sequential(nn.Linear((4 if edge_dim else 3) * Fin, Fin))])
        self.post_nns = nn.ModuleList([nn.Sequential(
            nn.Linear((len(x_aggregators) * len(self.scalers) + 1) * Fin, Fin))])
        self.rbf_emb = nn.Sequential(nn.Linear(num_radial, Fin), nn.Tanh())
        if edge_dim is not None:
            self.edge_encoder = nn.Linear(edge_dim, Fin)
        self.rbf_lin = nn.Linear(num_radial, Fin * 3, bias=False)
        self.scalar_message_mlp = nn.Sequential(nn.Linear(Fin, Fin), nn.Tanh(), nn.Linear(Fin, Fin), nn.SiLU(),
                                                nn.Linear(Fin, Fin * 3))

    _scale = _PNAScalers._scale

    def forward(self, x, v, edge_index, edge_rbf, edge_vec, edge_attr=None):
        Fin = self.node_size
        topo = getattr(edge_index, "_hgk_topo", None)
        if topo is None:   # `edge_index` is [E,2]: column 0 = i (aggregation side), column 1 = j
            topo = topo_of(edge_index.t().contiguous(), x.size(0))
        pre = self.pre_nns[0][0]
        w = pre.weight
        p = F.linear(x, w[:, :Fin], pre.bias)                 # x[src] block (+ bias)
        q = F.linear(x, w[:, Fin:2 * Fin])                    # x[dst] block
        m0 = edge_gather_add(p, q, topo, 0, 1) + F.linear(self.rbf_emb(edge_rbf), w[:, 2 * Fin:3 * Fin])
        if edge_attr is not None and self.edge_dim is not None:
            w_e = w[:, 3 * Fin:]
            m0 = m0 + edge_linear(edge_attr, w_e @ self.edge_encoder.weight, w_e @ self.edge_encoder.bias)
        f = self.scalar_message_mlp(m0) * self.rbf_lin(edge_rbf)
        g_sv, g_ev, m_s = f[:, :Fin], f[:, Fin:2 * Fin], f[:, 2 * Fin:]

        aggr = pna_aggregate(None, None, m_s, topo, side=0)   # [N, 4F] at i = edge_index[0]
        deg = topo.degree(0).to(x.dtype).view(-1, 1)
        delta_x = self.post_nns[0](torch.cat([x] + self._scale(aggr, deg), dim=-1))

        delta_v = contract_node(topo, 0, W=g_ev, U=edge_vec, C=3, CO=3, F=Fin)
        if v.requires_grad or bool(getattr(v, "_hgk_nonzero", True)):
            delta_v = delta_v + contract_node(topo, 0, A=v, side_a=1, W=g_sv, C=3, CO=3, F=Fin)
        return x + delta_x, v + delta_v


class PainnUpdate(nn.Module):
    """Node-level update (GEMM-bound, stays in torch)."""

    def __init__(self, node_size: int, last_layer=False):
        super().__init__()
        self.update_X = nn.Linear(node_size, node_size)
        self.update_V = nn.Linear(node_size, node_size)
        self.last_layer = last_layer
        self.update_mlp = nn.Sequential(nn.Linear(node_size * 2, node_size), nn.SiLU(),
                                        nn.Linear(node_size, node_size * (2 if last_layer else 3)))

    def forward(self, x, v):
        Xv = self.update_X(v)
        Vv = self.update_V(v)
        gates = self.update_mlp(torch.cat((torch.linalg.norm(Vv, dim=1), x), dim=-1))
        inner = torch.sum(Xv * Vv, dim=1)
        width = v.shape[-1]
        if self.last_layer:
            a_xv, a_xx = torch.split(gates, width, dim=-1)
            return x + a_xv * inner + a_xx
        a_vv, a_xv, a_xx = torch.split(gates, width, dim=-1)
        return x + a_xv * inner + a_xx, v + a_vv.unsqueeze(1) * Xv


class PNAEqStack(Base):
    def __init__(self, input_args, conv_args, deg: list, edge_dim: int, num_radial: int, radius: float,
                 *args, **kwargs):
        self.x_aggregators = ["mean", "min", "max", "std"]
        self.x_scalers = ["identity", "amplification", "attenuation", "linear", "inverse_linear"]
        self.deg = torch.Tensor(deg)
        self.edge_dim = edge_dim
        self.num_radial = num_radial
        self.radius = radius
        self.is_edge_model = True
        super().__init__(input_args, conv_args, *args, **kwargs)
        self.rbf = rbf_BasisLayer(self.num_radial, self.radius)

    def _init_conv(self):
        widths = [self.embed_dim] + [self.hidden_dim] * self.num_conv_layers
        for k, (w_in, w_out) in enumerate(zip(widths[:-1], widths[1:])):
            conv = self.get_conv(w_in, w_out, k == self.num_conv_layers - 1, edge_dim=self.edge_embed_dim)
            self.graph_convs.append(self._apply_global_attn(conv))
            self.feature_layers.append(nn.Identity())

    def get_conv(self, input_dim, output_dim, last_layer=False, edge_dim=None):
        width = output_dim if input_dim == 1 else input_dim
        assert width > 1, "PNAEq requires more than one hidden dimension between input_dim and output_dim."
        message = PainnMessage(node_size=input_dim, x_aggregators=self.x_aggregators, x_scalers=self.x_scalers,
                               deg=self.deg, edge_dim=edge_dim, num_radial=self.num_radial, pre_layers=1,
                               post_layers=1, divide_input=False)
        update = PainnUpdate(node_size=input_dim, last_layer=last_layer)
        node_embed_out = nn.Sequential(nn.Linear(input_dim, output_dim), nn.Tanh(),
                                       nn.Linear(output_dim, output_dim))
        steps = [(message, self.conv_args + " -> inv_node_feat, equiv_node_feat")]
        if last_layer:
            steps += [(update, "inv_node_feat, equiv_node_feat -> inv_node_feat"),
                      (node_embed_out, "inv_node_feat -> inv_node_feat")]
        else:
            steps += [(update, _BOTH), (node_embed_out, "inv_node_feat -> inv_node_feat"),
                      (nn.Linear(input_dim, output_dim), "equiv_node_feat -> equiv_node_feat")]
        steps.append((_passthrough, _BOTH))
        return Sequential(self.input_args, steps)

    def _embedding(self, data):
        super()._embedding(data)
        assert data.pos is not None, "PNAEq requires node positions (data.pos) to be set."
        unit, dist = get_edge_vectors_and_lengths(data.pos, data.edge_index, data.edge_shifts, normalize=True)
        edge = data.edge_index.t()
        edge._hgk_topo = topo_of(data.edge_index, data.pos.size(0))
        conv_args = {"edge_index": edge, "edge_rbf": self.rbf(dist.squeeze(-1)), "edge_vec": unit}
        if self.use_edge_attr:
            assert data.edge_attr is not None, "Data must have edge attributes if use_edge_attributes is set."
            conv_args["edge_attr"] = data.edge_attr
        x = self._gps_embeddings(data, conv_args) if self.use_global_attn else data.x
        v = torch.zeros(x.size(0), 3, x.size(1), device=x.device)
        v._hgk_nonzero = False   # first layer: the v * g_sv term of the message is identically zero
        return x, v, conv_args

    def __str__(self):
        return "PNAEqStack"
