"""Small torch modules the stacks are built from.

They keep the parameter / child names PyG gives them in the reference
(Sequential children `module_{k}`, BatchNorm's inner `module`), so a state dict
saved by the reference (hydragnn/utils/model/model.py:63-106) loads here and
vice versa.
"""
import math

import torch
from torch import nn


class Sequential(nn.Module):
    """Signature-driven sequential, `Sequential("a, b", [(fn, "a, b -> c"), ...])`,
    the container every reference get_conv returns (e.g. SCFStack.py:106-153)."""

    def __init__(self, input_args, modules):
        super().__init__()
        self._in = [a.strip() for a in input_args.split(",") if a.strip()]
        self._steps = []
        for idx, (fn, desc) in enumerate(modules):
            ins, outs = desc.split("->")
            ins = [a.strip() for a in ins.split(",") if a.strip()]
            outs = [a.strip() for a in outs.split(",") if a.strip()]
            if isinstance(fn, nn.Module):
                setattr(self, f"module_{idx}", fn)
                self._steps.append((f"module_{idx}", None, ins, outs))
            else:
                self._steps.append((None, fn, ins, outs))

    def forward(self, *args, **kwargs):
        env = dict(zip(self._in, args))
        for k, v in kwargs.items():
            if k not in self._in:
                raise TypeError(f"forward() got an unexpected keyword argument '{k}'")
            env[k] = v
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


class Linear(nn.Linear):
    """torch.nn.Linear (same parameters and state-dict keys); inside a CUDA-graph capture the bias
    gradient comes from hgk_colsum (ops/linear.py)."""

    def forward(self, x):
        from .ops import linear_rows

        return linear_rows(x, self.weight, self.bias)


class BatchNorm(nn.Module):
    """PyG BatchNorm: nn.BatchNorm1d under the child name `module`."""

    def __init__(self, in_channels, eps=1e-5, momentum=0.1, affine=True, track_running_stats=True):
        super().__init__()
        self.module = nn.BatchNorm1d(in_channels, eps, momentum, affine, track_running_stats)

    def reset_parameters(self):
        self.module.reset_parameters()

    def forward(self, x):
        return self.module(x)


class ShiftedSoftplus(nn.Module):
    def __init__(self):
        super().__init__()
        self.shift = math.log(2.0)

    def forward(self, x):
        return torch.nn.functional.softplus(x) - self.shift


class GaussianSmearing(nn.Module):
    """exp(coeff * (d - mu_k)^2), mu = linspace(start, stop, G), coeff = -0.5 / (mu_1 - mu_0)^2."""

    def __init__(self, start=0.0, stop=5.0, num_gaussians=50):
        super().__init__()
        offset = torch.linspace(start, stop, num_gaussians)
        self.coeff = -0.5 / (offset[1] - offset[0]).item() ** 2
        self.register_buffer("offset", offset)

    lazy = False   # SCFStack sets this: hand the distances on and let the consumer expand them

    def forward(self, dist):
        if self.lazy:
            return LazyGaussians(dist, self.offset, self.coeff)
        return LazyGaussians(dist, self.offset, self.coeff).materialize()


class LazyGaussians:
    """exp(coeff (d - mu_k)^2) not yet evaluated: the fused filter kernel expands the distances in
    registers, so the [E, G] tensor is only built if a consumer asks for it."""

    def __init__(self, dist, offset, coeff):
        self.dist, self.offset, self.coeff = dist, offset, coeff

    def size(self, dim=None):
        shape = (self.dist.numel(), self.offset.numel())
        return shape if dim is None else shape[dim]

    def materialize(self):
        delta = self.dist.view(-1, 1) - self.offset.view(1, -1)
        return torch.exp(self.coeff * torch.pow(delta, 2))


class FactoredEdgeAttr:
    """A per-edge embedding kept in factored form:  e = raw @ matrix.T  (raw [E, r], matrix [C, r]).

    With GPS the reference materialises e = rel_pos_emb(rel_pe) (or edge_lin(cat(edge_emb(edge_attr),
    rel_pos_emb(rel_pe)))), an [E, hidden] tensor, and every conv layer then feeds it to another
    Linear (hydragnn/models/Base.py:237-242 -> SCFStack.py:269-272, EGCLStack.py:240-245,
    PAINNStack.py:233-235).  All of those maps are linear and bias free, so a consumer Linear W
    applied to e equals raw @ (W @ matrix).T: the [E, hidden] tensor is never built, the E-row GEMM
    shrinks from K = hidden to K = r (8..9), and no gradient GEMM w.r.t. e is needed (raw is data).
    """

    def __init__(self, raw, matrix):
        self.raw, self.matrix = raw, matrix

    def size(self, dim=None):
        shape = (self.raw.size(0), self.matrix.size(0))
        return shape if dim is None else shape[dim]

    @property
    def shape(self):
        return self.size()

    def project(self, weight, bias=None):
        """F.linear(e, weight, bias) without forming e."""
        return torch.nn.functional.linear(self.raw, weight @ self.matrix, bias)

    def materialize(self):
        return torch.nn.functional.linear(self.raw, self.matrix)


def edge_linear(edge_attr, weight, bias=None):
    """F.linear on a plain or factored edge attribute."""
    if isinstance(edge_attr, FactoredEdgeAttr):
        return edge_attr.project(weight, bias)
    return torch.nn.functional.linear(edge_attr, weight, bias)


_ACTS = {"relu": nn.ReLU, "elu": nn.ELU, "gelu": nn.GELU, "silu": nn.SiLU, "tanh": nn.Tanh,
         "leaky_relu": nn.LeakyReLU, "sigmoid": nn.Sigmoid}


def activation_resolver(query="relu", **kwargs):
    if isinstance(query, nn.Module):
        return query
    return _ACTS[str(query).lower()](**kwargs)


def normalization_resolver(query, *args, **kwargs):
    if query is None:
        return None
    q = str(query).lower().replace("_", "")
    if q in ("batchnorm", "batch"):
        return BatchNorm(*args, **kwargs)
    if q in ("layernorm", "layer"):
        return nn.LayerNorm(*args, **kwargs)
    raise ValueError(f"normalization {query} is not supported")
