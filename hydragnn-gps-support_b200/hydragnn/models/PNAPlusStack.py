"""PNA+ stack on the B200 path (PNA with a DimeNet radial basis).

Same constructor / `get_conv` / `_embedding` contract and parameter names as the reference's
PNAPlusStack and its in-file PNAConv (hydragnn/models/PNAPlusStack.py:40-283).  The message

    h_e = pre([x_i, x_j, enc_e]) * rbf_lin(rbf_e),   enc_e = edge_encoder([edge_attr_e, rbf_emb(rbf_e)])
                                                          (or rbf_emb(rbf_e) without edge attributes)

is evaluated in split-weight form: the first two column blocks of `pre` act on node rows (two
N-row GEMMs), the third on the E-row encoding, and the four aggregations of the Hadamard product
run in one pass of the CSR kernel (ops.pna_aggregate with its per-edge multiplier): neither the
concatenation [E, 3F] nor the message tensor and its four scatter passes are materialised.
"""
import math

import torch
import torch.nn.functional as F
from torch import nn

from ..graph import topo_of
from ..nn import Sequential, edge_linear
from ..ops import pna_aggregate
from ..utils.model import get_edge_vectors_and_lengths
from .Base import Base
from .PNAStack import PNAConv as _PNAConvBase


def _passthrough(inv_node_feat, equiv_node_feat):
    return [inv_node_feat, equiv_node_feat]


class Envelope(nn.Module):
    """Polynomial cutoff of PyG's DimeNet (torch_geometric/nn/models/dimenet.py), x = d / cutoff."""

    def __init__(self, exponent: int):
        super().__init__()
        self.p = exponent + 1
        self.a = -(self.p + 1) * (self.p + 2) / 2
        self.b = self.p * (self.p + 2)
        self.c = -self.p * (self.p + 1) / 2

    def forward(self, x):
        x0 = x.pow(self.p - 1)
        x1 = x0 * x
        x2 = x1 * x
        return (1.0 / x + self.a * x0 + self.b * x1 + self.c * x2) * (x < 1.0).to(x.dtype)


class BesselBasisLayer(nn.Module):
    """envelope(d / cutoff) * sin(freq * d / cutoff); `freq` is a parameter initialised to pi * n."""

    def __init__(self, num_radial: int, cutoff: float = 5.0, envelope_exponent: int = 5):
        super().__init__()
        self.cutoff = cutoff
        self.envelope = Envelope(envelope_exponent)
        self.freq = nn.Parameter(torch.arange(1, num_radial + 1, dtype=torch.float32) * math.pi)

    def forward(self, dist):
        dist = dist.unsqueeze(-1) / self.cutoff
        return self.envelope(dist) * (self.freq * dist).sin()


class PNAConv(_PNAConvBase):
    """The reference's PNA+ convolution (PNAPlusStack.py:144-283), single tower."""

    def __init__(self, in_channels, out_channels, aggregators, scalers, deg, edge_dim=None, num_radial=5,
                 towers=1, pre_layers=1, post_layers=1, divide_input=False, **kwargs):
        super().__init__(in_channels, out_channels, aggregators, scalers, deg, edge_dim=None, towers=towers,
                         pre_layers=pre_layers, post_layers=post_layers, divide_input=divide_input)
        Fin = in_channels
        self.edge_dim = edge_dim
        self.pre_nns = nn.ModuleList([nn.Sequential(nn.Linear(3 * Fin, Fin))])     # always three blocks here
        self.rbf_lin = nn.Linear(num_radial, Fin, bias=False)
        self.rbf_emb = nn.Sequential(nn.Linear(num_radial, Fin), nn.ReLU())
        if edge_dim is not None:
            self.edge_encoder = nn.Linear(Fin + edge_dim, Fin)

    def forward(self, x, edge_index, rbf=None, edge_attr=None):
        topo = topo_of(edge_index, x.size(0))
        Fin = self.in_channels
        pre = self.pre_nns[0][0]
        p = F.linear(x, pre.weight[:, :Fin], pre.bias)       # x_i = x[dst] term (+ bias)
        q = F.linear(x, pre.weight[:, Fin:2 * Fin])          # x_j = x[src] term
        enc = self.rbf_emb(rbf)
        if edge_attr is not None and self.edge_dim is not None:
            w = self.edge_encoder.weight                     # [F, edge_dim + F] on cat[edge_attr, rbf_emb]
            de = w.size(1) - Fin
            enc = edge_linear(edge_attr, w[:, :de], self.edge_encoder.bias) + F.linear(enc, w[:, de:])
        r = F.linear(enc, pre.weight[:, 2 * Fin:])
        aggr = pna_aggregate(p, q, r, topo, M=self.rbf_lin(rbf))   # [N, 4F]: mean | min | max | std
        deg = topo.degree(1).to(x.dtype).view(-1, 1)
        out = torch.cat([x] + self._scale(aggr, deg), dim=-1)
        return self.lin(self.post_nns[0](out))


class PNAPlusStack(Base):
    def __init__(self, input_args, conv_args, deg: list, edge_dim: int, envelope_exponent: int,
                 num_radial: int, radius: float, *args, **kwargs):
        self.aggregators = ["mean", "min", "max", "std"]
        self.scalers = ["identity", "amplification", "attenuation", "linear"]
        self.deg = torch.Tensor(deg)
        self.edge_dim = edge_dim
        self.envelope_exponent = envelope_exponent
        self.num_radial = num_radial
        self.radius = radius
        self.is_edge_model = True
        super().__init__(input_args, conv_args, *args, **kwargs)
        self.rbf = BesselBasisLayer(self.num_radial, self.radius, self.envelope_exponent)

    def get_conv(self, input_dim, output_dim, edge_dim=None):
        pna = PNAConv(in_channels=input_dim, out_channels=output_dim, aggregators=self.aggregators,
                      scalers=self.scalers, deg=self.deg, edge_dim=edge_dim, num_radial=self.num_radial,
                      pre_layers=1, post_layers=1, divide_input=False)
        return Sequential(self.input_args, [
            (pna, self.conv_args + " -> inv_node_feat"),
            (_passthrough, "inv_node_feat, equiv_node_feat -> inv_node_feat, equiv_node_feat")])

    def _embedding(self, data):
        super()._embedding(data)
        assert data.pos is not None, "PNA+ requires node positions (data.pos) to be set."
        _, dist = get_edge_vectors_and_lengths(data.pos, data.edge_index, data.edge_shifts)
        conv_args = {"edge_index": data.edge_index, "rbf": self.rbf(dist.squeeze(-1))}
        if self.use_edge_attr:
            assert data.edge_attr is not None, "Data must have edge attributes if use_edge_attributes is set."
            conv_args["edge_attr"] = data.edge_attr
        if self.use_global_attn:
            return self._gps_embeddings(data, conv_args), data.pos, conv_args
        return data.x, data.pos, conv_args

    def __str__(self):
        return "PNAStack"
