"""PNA stack on the B200 path.

Same constructor / `get_conv` contract and parameter names as the reference's PNAStack,
which wraps PyG's PNAConv(towers=1, pre_layers=1, post_layers=1, divide_input=False) with
aggregators mean/min/max/std and scalers identity/amplification/attenuation/linear
(hydragnn/models/PNAStack.py:19-70).  Underneath, the message Linear on
cat[x_i, x_j(, enc(e))] is split into node-level GEMMs and the four aggregations run in
one pass of a CSR kernel (ops.pna_aggregate) that forms each message in registers; the
[E, F] message tensor and its four scatter passes are never materialised.
"""
import torch
import torch.nn.functional as F
from torch import nn

from ..graph import topo_of
from ..nn import Sequential, edge_linear
from ..ops import pna_aggregate
from .Base import Base


def _passthrough(inv_node_feat, equiv_node_feat):
    return [inv_node_feat, equiv_node_feat]


class PNAConv(nn.Module):
    """Principal neighbourhood aggregation, single tower, one pre / post Linear."""

    def __init__(self, in_channels, out_channels, aggregators, scalers, deg, edge_dim=None,
                 towers=1, pre_layers=1, post_layers=1, divide_input=False, **kwargs):
        super().__init__()
        if towers != 1 or pre_layers != 1 or post_layers != 1 or divide_input:
            raise NotImplementedError("only the reference's PNAConv configuration is on the B200 path")
        if list(aggregators) != ["mean", "min", "max", "std"]:
            raise NotImplementedError(f"aggregators {aggregators}: the kernel computes mean,min,max,std")
        self.in_channels, self.out_channels, self.edge_dim = in_channels, out_channels, edge_dim
        self.scalers = list(scalers)
        hist = torch.as_tensor(deg, dtype=torch.float)
        bins = torch.arange(hist.numel(), dtype=torch.float)
        total = int(hist.sum())
        self.avg_deg_lin = float((bins * hist).sum()) / total
        self.avg_deg_log = float(((bins + 1).log() * hist).sum()) / total
        Fin = in_channels
        if edge_dim is not None:
            self.edge_encoder = nn.Linear(edge_dim, Fin)
        self.pre_nns = nn.ModuleList([nn.Sequential(nn.Linear((3 if edge_dim else 2) * Fin, Fin))])
        width = (len(aggregators) * len(self.scalers) + 1) * Fin
        self.post_nns = nn.ModuleList([nn.Sequential(nn.Linear(width, out_channels))])
        self.lin = nn.Linear(out_channels, out_channels)

    def _scale(self, aggr, deg):
        outs = []
        for s in self.scalers:
            if s == "identity":
                outs.append(aggr)
            elif s == "amplification":
                outs.append(aggr * (torch.log(deg + 1) / self.avg_deg_log))
            elif s == "attenuation":
                outs.append(aggr * (self.avg_deg_log / torch.log(deg.clamp(min=1) + 1)))
            elif s == "linear":
                outs.append(aggr * (deg / self.avg_deg_lin))
            elif s == "inverse_linear":
                outs.append(aggr * (self.avg_deg_lin / deg.clamp(min=1)))
            else:
                raise ValueError(f"Unknown scaler {s}")
        return outs

    def forward(self, x, edge_index, edge_attr=None):
        topo = topo_of(edge_index, x.size(0))
        Fin = self.in_channels
        pre = self.pre_nns[0][0]
        p = F.linear(x, pre.weight[:, :Fin], pre.bias)       # x_i = x[dst] term (+ bias)
        q = F.linear(x, pre.weight[:, Fin:2 * Fin])          # x_j = x[src] term
        r = None
        if edge_attr is not None and self.edge_dim is not None:
            # pre[:, 2F:] (edge_encoder(e)) folded into one map of width edge_dim
            w_e = pre.weight[:, 2 * Fin:]
            r = edge_linear(edge_attr, w_e @ self.edge_encoder.weight, w_e @ self.edge_encoder.bias)
        aggr = pna_aggregate(p, q, r, topo)                  # [N, 4F]: mean | min | max | std
        deg = topo.degree(1).to(x.dtype).view(-1, 1)
        out = torch.cat([x] + self._scale(aggr, deg), dim=-1)
        return self.lin(self.post_nns[0](out))


class PNAStack(Base):
    def __init__(self, input_args, conv_args, deg: list, edge_dim: int, *args, **kwargs):
        self.aggregators = ["mean", "min", "max", "std"]
        self.scalers = ["identity", "amplification", "attenuation", "linear"]
        self.deg = torch.Tensor(deg)
        self.edge_dim = edge_dim
        self.is_edge_model = True
        super().__init__(input_args, conv_args, *args, **kwargs)

    def get_conv(self, input_dim, output_dim, edge_dim=None):
        pna = PNAConv(in_channels=input_dim, out_channels=output_dim, aggregators=self.aggregators,
                      scalers=self.scalers, deg=self.deg, edge_dim=edge_dim, pre_layers=1,
                      post_layers=1, divide_input=False)
        return Sequential(self.input_args, [
            (pna, self.conv_args + " -> inv_node_feat"),
            (_passthrough, "inv_node_feat, equiv_node_feat -> inv_node_feat, equiv_node_feat")])

    def __str__(self):
        return "PNAStack"
