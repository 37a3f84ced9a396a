"""ORACLE (test infrastructure only -- never imported by the product path).

CPU restatement, in plain torch/numpy, of the THIRD-PARTY library semantics the
reference's hot path relies on.  None of these libraries is vendored under
/root/reference or installed in this image, so their published behaviour is
restated here, each piece citing the reference call site that depends on it:

  torch_geometric == 2.5.2 (requirements-pyg.txt:1)
      nn.Sequential             hydragnn/models/*Stack.py get_conv
      nn.MessagePassing(add)    hydragnn/models/SCFStack.py:214,285
      nn.PNAConv + DegreeScalerAggregation   hydragnn/models/PNAStack.py:43-53
      nn.models.dimenet.BesselBasisLayer     hydragnn/models/PNAPlusStack.py:32,73
      nn.BatchNorm, global_mean_pool         hydragnn/models/Base.py:15,206,478
      nn.models.schnet.{GaussianSmearing, ShiftedSoftplus, RadiusInteractionGraph}
                                hydragnn/models/SCFStack.py:21-25,58-61
      utils.to_dense_batch      hydragnn/globalAtt/gps.py:126
      transforms.RadiusGraph / Distance      preprocess/graph_samples_checks_and_updates.py:109-114
  torch_scatter == 2.1.2        hydragnn/models/Base.py:599 (scatter_add)
  torch_cluster == 1.6.3        radius_graph behind RadiusGraph / RadiusInteractionGraph

Parity status of THIS file: "parity unpinned" -- the libraries cannot be run
here; the semantics are recalled from the pinned versions and
cross-checked only against the reference's in-repo near-copy of PNAConv
(hydragnn/models/PNAPlusStack.py:144-283) and its neighbour-count known answers.
What IS pinned (tests/golden) is the reference's own module code running on
top of this layer.
"""
import inspect
import math
from typing import List, Optional

import numpy as np
import torch
from torch import Tensor
from torch.nn import Linear, ModuleList


# --------------------------------------------------------------- scatter ---
def scatter_add(src: Tensor, index: Tensor, dim: int = 0, dim_size: Optional[int] = None) -> Tensor:
    """torch_scatter.scatter_add along dim 0 (sequential index_add_ on CPU)."""
    assert dim == 0
    if dim_size is None:
        dim_size = int(index.max()) + 1 if index.numel() > 0 else 0
    out = src.new_zeros((dim_size,) + tuple(src.shape[1:]))
    return out.index_add_(0, index, src)


def scatter(src: Tensor, index: Tensor, dim: int = 0, dim_size: Optional[int] = None,
            reduce: str = "sum") -> Tensor:
    """torch_geometric.utils.scatter for sum / mean / min / max (PyG 2.5.2 semantics:
    mean = sum / clamp(count, 1); min/max use scatter_reduce(amin/amax, include_self=False)
    on a zero tensor, so empty segments stay 0)."""
    assert dim == 0
    if dim_size is None:
        dim_size = int(index.max()) + 1 if index.numel() > 0 else 0
    if reduce in ("sum", "add"):
        return scatter_add(src, index, 0, dim_size)
    if reduce == "mean":
        total = scatter_add(src, index, 0, dim_size)
        count = src.new_zeros(dim_size).index_add_(0, index, src.new_ones(index.numel()))
        count = count.clamp(min=1).view((-1,) + (1,) * (src.dim() - 1))
        return total / count
    if reduce in ("min", "max"):
        idx = index.view((-1,) + (1,) * (src.dim() - 1)).expand_as(src)
        out = src.new_zeros((dim_size,) + tuple(src.shape[1:]))
        return out.scatter_reduce_(0, idx, src, reduce="a" + reduce, include_self=False)
    raise ValueError(reduce)


def global_mean_pool(x: Tensor, batch: Optional[Tensor], size: Optional[int] = None) -> Tensor:
    if batch is None:
        return x.mean(dim=0, keepdim=True)
    if size is None:
        size = int(batch.max()) + 1
    return scatter(x, batch, 0, size, "mean")


def degree(index: Tensor, num_nodes: Optional[int] = None, dtype=None) -> Tensor:
    n = int(index.max()) + 1 if num_nodes is None else num_nodes
    out = torch.zeros(n, dtype=dtype or torch.get_default_dtype())
    return out.index_add_(0, index, torch.ones(index.numel(), dtype=out.dtype))


def to_dense_batch(x: Tensor, batch: Optional[Tensor] = None, fill_value: float = 0.0,
                   max_num_nodes: Optional[int] = None, batch_size: Optional[int] = None):
    """PyG to_dense_batch: batch=None -> (x[None], all-True mask [1,N])."""
    if batch is None and max_num_nodes is None:
        mask = torch.ones(1, x.size(0), dtype=torch.bool, device=x.device)
        return x.unsqueeze(0), mask
    if batch is None:
        batch = x.new_zeros(x.size(0), dtype=torch.long)
    if batch_size is None:
        batch_size = int(batch.max()) + 1
    num_nodes = torch.zeros(batch_size, dtype=torch.long).index_add_(
        0, batch, torch.ones_like(batch))
    cum = torch.cat([num_nodes.new_zeros(1), num_nodes.cumsum(0)])
    if max_num_nodes is None:
        max_num_nodes = int(num_nodes.max())
    tmp = torch.arange(batch.size(0)) - cum[batch]
    idx = tmp + batch * max_num_nodes
    out = x.new_full((batch_size * max_num_nodes,) + tuple(x.shape[1:]), fill_value)
    out[idx] = x
    out = out.view((batch_size, max_num_nodes) + tuple(x.shape[1:]))
    mask = torch.zeros(batch_size * max_num_nodes, dtype=torch.bool)
    mask[idx] = True
    return out, mask.view(batch_size, max_num_nodes)


# ---------------------------------------------------------------- modules ---
class BatchNorm(torch.nn.Module):
    """torch_geometric.nn.BatchNorm = wrapper around nn.BatchNorm1d (state-dict key `module.*`)."""

    def __init__(self, in_channels, eps=1e-5, momentum=0.1, affine=True, track_running_stats=True):
        super().__init__()
        self.module = torch.nn.BatchNorm1d(in_channels, eps, momentum, affine, track_running_stats)

    def reset_parameters(self):
        self.module.reset_parameters()

    def forward(self, x):
        return self.module(x)


class Sequential(torch.nn.Module):
    """torch_geometric.nn.Sequential('a, b', [(fn, 'a, b -> c'), ...]).

    Children that are nn.Modules are registered as module_{position}."""

    def __init__(self, input_args: str, modules: list):
        super().__init__()
        self._in = [a.strip() for a in input_args.split(",") if a.strip()]
        self._steps = []
        for idx, entry in enumerate(modules):
            fn, desc = entry
            ins, outs = desc.split("->")
            ins = [a.strip() for a in ins.split(",") if a.strip()]
            outs = [a.strip() for a in outs.split(",") if a.strip()]
            name = f"module_{idx}"
            if isinstance(fn, torch.nn.Module):
                setattr(self, name, fn)
                self._steps.append((name, None, ins, outs))
            else:
                self._steps.append((None, fn, ins, outs))

    def reset_parameters(self):
        for m in self.children():
            if hasattr(m, "reset_parameters"):
                m.reset_parameters()

    def forward(self, *args, **kwargs):
        env = dict(zip(self._in, args))
        for k, v in kwargs.items():
            if k not in self._in:
                raise TypeError(f"forward() got an unexpected keyword argument '{k}'")
            env[k] = v
        missing = [a for a in self._in if a not in env]
        if missing:
            raise TypeError(f"forward() missing arguments {missing}")
        out = None
        for name, fn, ins, outs in self._steps:
            f = getattr(self, name) if name is not None else fn
            out = f(*[env[a] for a in ins])
            if len(outs) == 1:
                env[outs[0]] = out
            else:
                for o, val in zip(outs, out):
                    env[o] = val
                out = tuple(env[o] for o in outs)
        return out


class MessagePassing(torch.nn.Module):
    """Message passing, flow source_to_target: x_j = x[edge_index[0]], x_i = x[edge_index[1]],
    aggregation at edge_index[1] -- a sum, or an aggregation module called as
    aggr(messages, index, dim_size) (PyG's `aggr_module`, e.g. DegreeScalerAggregation)."""

    def __init__(self, aggr="add", flow="source_to_target", node_dim=0, **kwargs):
        super().__init__()
        assert flow == "source_to_target" and node_dim == 0
        if isinstance(aggr, torch.nn.Module):
            self.aggr_module = aggr
            self.aggr = None
        else:
            assert aggr in ("add", "sum")
            self.aggr_module = None
            self.aggr = aggr

    def reset_parameters(self):
        pass

    def propagate(self, edge_index, size=None, **kwargs):
        params = list(inspect.signature(self.message).parameters)
        num_nodes = None
        for v in kwargs.values():
            if isinstance(v, Tensor) and v.dim() >= 1:
                num_nodes = v.size(0)
                break
        args = []
        for p in params:
            if p.endswith("_j"):
                base = kwargs[p[:-2]]
                num_nodes = base.size(0)
                args.append(base.index_select(0, edge_index[0]))
            elif p.endswith("_i"):
                base = kwargs[p[:-2]]
                num_nodes = base.size(0)
                args.append(base.index_select(0, edge_index[1]))
            else:
                args.append(kwargs.get(p))
        msg = self.message(*args)
        if self.aggr_module is not None:
            return self.aggr_module(msg, edge_index[1], num_nodes)
        return scatter_add(msg, edge_index[1], 0, num_nodes)


class ShiftedSoftplus(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.shift = torch.log(torch.tensor(2.0)).item()

    def forward(self, x):
        return torch.nn.functional.softplus(x) - self.shift


class GaussianSmearing(torch.nn.Module):
    def __init__(self, start=0.0, stop=5.0, num_gaussians=50):
        super().__init__()
        offset = torch.linspace(start, stop, num_gaussians)
        self.coeff = -0.5 / (offset[1] - offset[0]).item() ** 2
        self.register_buffer("offset", offset)

    def forward(self, dist):
        dist = dist.view(-1, 1) - self.offset.view(1, -1)
        return torch.exp(self.coeff * torch.pow(dist, 2))


# ------------------------------------------------------------- neighbours ---
def radius_graph_cluster(pos: Tensor, r: float, batch: Optional[Tensor] = None, loop: bool = False,
                         max_num_neighbors: int = 32) -> Tensor:
    """torch_cluster.radius_graph (CPU semantics): for every centre j, sources i of the same
    batch segment with |x_i-x_j|^2 < r^2 in fp32 (squares summed x,y,z; no FMA), nearest first,
    at most max_num_neighbors (self excluded unless loop); output rows [source; centre],
    grouped by centre.  Ties in distance are broken by source index (our rule)."""
    x = pos.detach().cpu().numpy().astype(np.float32)
    n = x.shape[0]
    b = np.zeros(n, dtype=np.int64) if batch is None else batch.detach().cpu().numpy()
    r2 = np.float32(float(r) * float(r))
    src_all, dst_all = [], []
    order = np.arange(n)
    for g in np.unique(b):
        ids = order[b == g]
        xs = x[ids]
        d = xs[:, None, :] - xs[None, :, :]           # d[i, j] = x_i - x_j
        sq = (d * d).astype(np.float32)
        d2 = ((sq[..., 0] + sq[..., 1]).astype(np.float32) + sq[..., 2]).astype(np.float32)
        for jj, j in enumerate(ids):
            m = d2[:, jj] < r2
            if not loop:
                m[jj] = False
            cand = np.nonzero(m)[0]
            key = np.lexsort((ids[cand], d2[cand, jj]))
            cand = cand[key][: max_num_neighbors if max_num_neighbors else None]
            src_all.append(ids[cand])
            dst_all.append(np.full(cand.size, j, dtype=np.int64))
    src = np.concatenate(src_all) if src_all else np.zeros(0, np.int64)
    dst = np.concatenate(dst_all) if dst_all else np.zeros(0, np.int64)
    return torch.from_numpy(np.stack([src, dst]).astype(np.int64))


class RadiusInteractionGraph(torch.nn.Module):
    """PyG schnet.RadiusInteractionGraph: radius_graph + edge_weight = |pos[row]-pos[col]|."""

    def __init__(self, cutoff=10.0, max_num_neighbors=32):
        super().__init__()
        self.cutoff, self.max_num_neighbors = cutoff, max_num_neighbors

    def forward(self, pos, batch):
        edge_index = radius_graph_cluster(pos, self.cutoff, batch, False, self.max_num_neighbors)
        row, col = edge_index
        edge_weight = (pos[row] - pos[col]).norm(dim=-1)
        return edge_index, edge_weight


def neighbor_list_ase(pos, cell, pbc, cutoff: float, self_interaction: bool = True):
    """ase.neighborlist.neighbor_list("ijdS", ...) semantics, brute force in fp64.

    All (i, j, S) with d = |pos[j] - pos[i] + S.cell| < cutoff (strict), i == j kept for S != 0
    and for S == 0 when self_interaction.  Positions are NOT wrapped (S refers to the positions
    as given).  Returned grouped by i then j then S, which is an implementation detail callers
    must not rely on (the reference re-sorts by (dst, length), :259)."""
    pos = np.asarray(pos, dtype=np.float64)
    cell = np.asarray(cell, dtype=np.float64).reshape(3, 3)
    pbc = np.asarray(pbc, dtype=bool).reshape(3)
    n = pos.shape[0]
    inv = np.linalg.inv(cell)
    heights = 1.0 / np.linalg.norm(inv, axis=0)
    frac = pos @ inv
    reps = []
    for k in range(3):
        if pbc[k]:
            span = frac[:, k].max() - frac[:, k].min() if n else 0.0
            reps.append(int(math.ceil(cutoff / heights[k] + span)))
        else:
            reps.append(0)
    out_i, out_j, out_d, out_S = [], [], [], []
    rng = [range(-m, m + 1) for m in reps]
    for sx in rng[0]:
        for sy in rng[1]:
            for sz in rng[2]:
                S = np.array([sx, sy, sz], dtype=np.float64)
                shift = S[0] * cell[0] + S[1] * cell[1] + S[2] * cell[2]
                D = (pos[None, :, :] - pos[:, None, :]) + shift[None, None, :]   # D[i, j]
                d = np.sqrt(D[..., 0] * D[..., 0] + D[..., 1] * D[..., 1] + D[..., 2] * D[..., 2])
                m = d < cutoff
                if sx == 0 and sy == 0 and sz == 0 and not self_interaction:
                    m[np.arange(n), np.arange(n)] = False
                ii, jj = np.nonzero(m)
                out_i.append(ii); out_j.append(jj); out_d.append(d[ii, jj])
                out_S.append(np.tile(np.array([sx, sy, sz], dtype=np.int64), (ii.size, 1)))
    i = np.concatenate(out_i); j = np.concatenate(out_j)
    d = np.concatenate(out_d); S = np.concatenate(out_S)
    order = np.lexsort((S[:, 2], S[:, 1], S[:, 0], j, i))
    return i[order], j[order], d[order], S[order]


# -------------------------------------------------------------------- PNA ---
class DegreeScalerAggregation(torch.nn.Module):
    """PyG DegreeScalerAggregation(aggr=[mean,min,max,std], scaler=[identity, amplification,
    attenuation, linear], deg histogram)."""

    def __init__(self, aggr: List[str], scaler: List[str], deg: Tensor, train_norm: bool = False):
        super().__init__()
        self.aggr, self.scaler = list(aggr), list(scaler)
        deg = deg.to(torch.float)
        num_nodes = int(deg.sum())
        bin_degrees = torch.arange(deg.numel(), dtype=torch.float)
        self.avg_deg = {
            "lin": float((bin_degrees * deg).sum()) / num_nodes,
            "log": float(((bin_degrees + 1).log() * deg).sum()) / num_nodes,
            "exp": float((bin_degrees.exp() * deg).sum()) / num_nodes,
        }

    def forward(self, x: Tensor, index: Tensor, dim_size: int) -> Tensor:
        outs = []
        for a in self.aggr:
            if a in ("mean", "min", "max", "sum"):
                outs.append(scatter(x, index, 0, dim_size, a))
            elif a == "std":
                mean = scatter(x, index, 0, dim_size, "mean")
                mean2 = scatter(x * x, index, 0, dim_size, "mean")
                var = mean2 - mean * mean
                out = var.clamp(min=1e-5).sqrt()
                outs.append(out.masked_fill(out <= math.sqrt(1e-5), 0.0))
            else:
                raise ValueError(a)
        out = torch.cat(outs, dim=-1)
        deg = degree(index, dim_size, dtype=out.dtype)
        size = [1] * out.dim()
        size[0] = -1
        deg = deg.view(size)
        scaled = []
        for s in self.scaler:
            if s == "identity":
                scaled.append(out)
            elif s == "amplification":
                scaled.append(out * (torch.log(deg + 1) / self.avg_deg["log"]))
            elif s == "attenuation":
                scaled.append(out * (self.avg_deg["log"] / torch.log(deg.clamp(min=1) + 1)))
            elif s == "linear":
                scaled.append(out * (deg / self.avg_deg["lin"]))
            elif s == "inverse_linear":
                scaled.append(out * (self.avg_deg["lin"] / deg.clamp(min=1)))
            else:
                raise ValueError(s)
        return torch.cat(scaled, dim=-1) if len(scaled) > 1 else scaled[0]


class PNAConv(torch.nn.Module):
    """PyG PNAConv (towers, pre/post layers of plain Linear when *_layers == 1)."""

    def __init__(self, in_channels, out_channels, aggregators, scalers, deg, edge_dim=None,
                 towers=1, pre_layers=1, post_layers=1, divide_input=False, act="relu",
                 act_kwargs=None, train_norm=False, **kwargs):
        super().__init__()
        assert pre_layers == 1 and post_layers == 1, "only the reference's configuration is restated"
        self.aggr_module = DegreeScalerAggregation(aggregators, scalers, deg, train_norm)
        self.in_channels, self.out_channels = in_channels, out_channels
        self.edge_dim, self.towers, self.divide_input = edge_dim, towers, divide_input
        self.F_in = in_channels // towers if divide_input else in_channels
        self.F_out = out_channels // towers
        if edge_dim is not None:
            self.edge_encoder = Linear(edge_dim, self.F_in)
        self.pre_nns = ModuleList()
        self.post_nns = ModuleList()
        for _ in range(towers):
            self.pre_nns.append(torch.nn.Sequential(Linear((3 if edge_dim else 2) * self.F_in, self.F_in)))
            width = (len(aggregators) * len(scalers) + 1) * self.F_in
            self.post_nns.append(torch.nn.Sequential(Linear(width, self.F_out)))
        self.lin = Linear(out_channels, out_channels)

    def reset_parameters(self):
        for m in self.modules():
            if isinstance(m, Linear):
                m.reset_parameters()

    def forward(self, x, edge_index, edge_attr=None):
        if self.divide_input:
            x = x.view(-1, self.towers, self.F_in)
        else:
            x = x.view(-1, 1, self.F_in).repeat(1, self.towers, 1)
        x_j = x.index_select(0, edge_index[0])
        x_i = x.index_select(0, edge_index[1])
        if edge_attr is not None:
            e = self.edge_encoder(edge_attr).view(-1, 1, self.F_in).repeat(1, self.towers, 1)
            h = torch.cat([x_i, x_j, e], dim=-1)
        else:
            h = torch.cat([x_i, x_j], dim=-1)
        hs = torch.stack([nn(h[:, i]) for i, nn in enumerate(self.pre_nns)], dim=1)
        out = self.aggr_module(hs, edge_index[1], x.size(0))
        out = torch.cat([x, out], dim=-1)
        outs = [nn(out[:, i]) for i, nn in enumerate(self.post_nns)]
        return self.lin(torch.cat(outs, dim=1))


# ------------------------------------------------------- DimeNet radial basis ---
class Envelope(torch.nn.Module):
    """torch_geometric.nn.models.dimenet.Envelope: smooth polynomial cutoff on x = d / cutoff."""

    def __init__(self, exponent: int):
        super().__init__()
        self.p = exponent + 1
        self.a = -(self.p + 1) * (self.p + 2) / 2
        self.b = self.p * (self.p + 2)
        self.c = -self.p * (self.p + 1) / 2

    def forward(self, x: Tensor) -> Tensor:
        p, a, b, c = self.p, self.a, self.b, self.c
        x_pow_p0 = x.pow(p - 1)
        x_pow_p1 = x_pow_p0 * x
        x_pow_p2 = x_pow_p1 * x
        return (1.0 / x + a * x_pow_p0 + b * x_pow_p1 + c * x_pow_p2) * (x < 1.0).to(x.dtype)


class BesselBasisLayer(torch.nn.Module):
    """torch_geometric.nn.models.dimenet.BesselBasisLayer (used by PNAPlusStack.py:73-75,114):
    envelope(d / cutoff) * sin(freq * d / cutoff), freq a parameter initialised to pi * (1..num_radial)."""

    def __init__(self, num_radial: int, cutoff: float = 5.0, envelope_exponent: int = 5):
        super().__init__()
        self.cutoff = cutoff
        self.envelope = Envelope(envelope_exponent)
        self.freq = torch.nn.Parameter(torch.empty(num_radial))
        self.reset_parameters()

    def reset_parameters(self):
        with torch.no_grad():
            torch.arange(1, self.freq.numel() + 1, out=self.freq).mul_(math.pi)
        self.freq.requires_grad_()

    def forward(self, dist: Tensor) -> Tensor:
        dist = dist.unsqueeze(-1) / self.cutoff
        return self.envelope(dist) * (self.freq * dist).sin()


# ----------------------------------------------------------- resolvers etc ---
def activation_resolver(query="relu", *args, **kwargs):
    if isinstance(query, torch.nn.Module):
        return query
    table = {"relu": torch.nn.ReLU, "elu": torch.nn.ELU, "gelu": torch.nn.GELU,
             "silu": torch.nn.SiLU, "tanh": torch.nn.Tanh, "leakyrelu": torch.nn.LeakyReLU,
             "leaky_relu": torch.nn.LeakyReLU, "sigmoid": torch.nn.Sigmoid}
    return table[str(query).lower().replace("_", "") if str(query).lower() not in table else str(query).lower()](*args, **kwargs)


def normalization_resolver(query, *args, **kwargs):
    if query is None:
        return None
    q = str(query).lower().replace("_", "")
    if q in ("batchnorm", "batch"):
        return BatchNorm(*args, **kwargs)
    if q in ("layernorm", "layer"):
        return torch.nn.LayerNorm(*args, **kwargs)
    raise ValueError(f"normalization {query} not restated")


def reset(value):
    if hasattr(value, "reset_parameters"):
        value.reset_parameters()
    else:
        for child in value.children() if hasattr(value, "children") else []:
            reset(child)


# -- torch_geometric.transforms.AddLaplacianEigenvectorPE (2.5.2) ---------------------------------------------
# reference call site: hydragnn/preprocess/serialized_dataset_loader.py:90-94,183 (k = pe_dim, is_undirected=True)
def laplacian_eigenvector_pe(edge_index: Tensor, num_nodes: int, k: int, signs: Optional[Tensor] = None):
    """pe [N, k]: eigenvectors 1..k (ascending eigenvalue; the trivial one skipped) of the symmetric
    normalised Laplacian  L = I - D^-1/2 A D^-1/2  built the PyG way: get_laplacian(normalization="sym")
    sums duplicate edges, takes the degree as the ROW sum of A, and zeroes 1/sqrt(deg) of isolated
    nodes; N < 100 -> numpy.linalg.eigh of the dense matrix, otherwise scipy eigsh(k + 1, which="SA")
    (is_undirected=True).  PyG then multiplies every column by a RANDOM sign (-1 + 2 randint(0, 2));
    `signs` ([k] of +-1) stands for that draw, None leaves the solver's signs."""
    import numpy as np

    N = int(num_nodes)
    row, col = edge_index[0].numpy(), edge_index[1].numpy()
    A = np.zeros((N, N), dtype=np.float64)
    np.add.at(A, (row, col), 1.0)
    deg = A.sum(1)
    with np.errstate(divide="ignore"):
        dis = 1.0 / np.sqrt(deg)
    dis[np.isinf(dis)] = 0.0
    L = np.eye(N) - dis[:, None] * A * dis[None, :]
    if N < 100:
        vals, vecs = np.linalg.eigh(L)
    else:
        from scipy.sparse import csr_matrix
        from scipy.sparse.linalg import eigsh

        vals, vecs = eigsh(csr_matrix(L), k=k + 1, which="SA", return_eigenvectors=True)
    vecs = np.real(vecs[:, vals.argsort()])
    pe = torch.from_numpy(vecs[:, 1:k + 1].copy()).to(torch.float32)
    if signs is not None:
        pe = pe * signs.to(pe.dtype)
    return pe
