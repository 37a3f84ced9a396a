"""Minimal Data / Batch containers (the PyG-free stand-in for torch_geometric.data).

Field contract: SURVEY.md Appendix B.  `Batch.from_data_list` concatenates node
and edge tensors with node offsets, builds `batch` / `ptr`, and -- unlike the
reference, which calls `.unique()` on the device every forward
(hydragnn/models/Base.py:484-485) -- keeps the graph sizes on the host so that no
device->host sync is needed later.
"""
import torch

from .ops.topo import Topo

_NODE_KEYS = ("x", "pos", "pe", "forces", "z")
_EDGE_KEYS = ("edge_attr", "edge_shifts", "rel_pe")
_GRAPH_KEYS = ("y", "energy", "cell", "pbc", "y_loc", "dataset_name", "graph_attr")


class Data:
    def __init__(self, **kwargs):
        for k, v in kwargs.items():
            setattr(self, k, v)

    def __contains__(self, key):
        return getattr(self, key, None) is not None

    def keys(self):
        return [k for k, v in self.__dict__.items() if not k.startswith("_") and v is not None]

    @property
    def num_nodes(self):
        for k in ("x", "pos"):
            v = getattr(self, k, None)
            if v is not None:
                return v.size(0)
        return 0

    @property
    def num_edges(self):
        ei = getattr(self, "edge_index", None)
        return 0 if ei is None else ei.size(1)

    def to(self, device, non_blocking=False):
        out = self.__class__.__new__(self.__class__)
        for k, v in self.__dict__.items():
            if torch.is_tensor(v):
                out.__dict__[k] = v.to(device, non_blocking=non_blocking)
            elif isinstance(v, Topo):
                out.__dict__[k] = None
            else:
                out.__dict__[k] = v
        return out

    def clone(self):
        out = self.__class__.__new__(self.__class__)
        for k, v in self.__dict__.items():
            out.__dict__[k] = v.clone() if torch.is_tensor(v) else v
        return out

    def pin_memory(self):
        for k, v in self.__dict__.items():
            if torch.is_tensor(v) and not v.is_cuda:
                self.__dict__[k] = v.pin_memory()
        return self


class Batch(Data):
    @classmethod
    def from_data_list(cls, data_list):
        out = cls()
        sizes = [d.num_nodes for d in data_list]
        ptr = [0]
        for n in sizes:
            ptr.append(ptr[-1] + n)
        out.num_graphs = len(data_list)
        out.ptr = torch.tensor(ptr, dtype=torch.long)
        out.batch = torch.repeat_interleave(torch.arange(len(data_list)), torch.tensor(sizes))
        out.graph_sizes = sizes                       # host copy: no .unique() later
        keys = set()
        for d in data_list:
            keys.update(d.keys())
        for k in sorted(keys):
            vals = [getattr(d, k, None) for d in data_list]
            if any(v is None for v in vals) or not torch.is_tensor(vals[0]):
                continue
            if k == "edge_index":
                out.edge_index = torch.cat([v + off for v, off in zip(vals, ptr[:-1])], dim=1)
            elif k in ("cell", "pbc"):
                out.__dict__[k] = torch.stack([v.reshape(3, 3) if k == "cell" else v.reshape(3) for v in vals])
            elif k in ("y_loc", "dataset_name"):
                out.__dict__[k] = torch.cat([v.reshape(1, -1) for v in vals], dim=0)
            elif k == "energy":
                out.energy = torch.cat([v.reshape(-1) for v in vals])
            else:
                out.__dict__[k] = torch.cat(vals, dim=0)
        return out

    def to(self, device, non_blocking=False):
        out = super().to(device, non_blocking)
        return out


def topo_of(edge_index, num_nodes, sorted_side=None):
    """Topo cached on the edge_index tensor object (built once per batch)."""
    topo = getattr(edge_index, "_hgk_topo", None)
    if topo is None or topo.num_rows[0] != num_nodes:
        topo = Topo.from_edge_index(edge_index, num_nodes, sorted_side=sorted_side)
        try:
            edge_index._hgk_topo = topo
        except Exception:
            pass
    return topo


def pool_topo(data, num_nodes):
    """One-sided topology node -> graph for pooling (batch is sorted by construction)."""
    topo = getattr(data, "_pool_topo", None)
    if topo is None:
        batch = data.batch
        ptr = getattr(data, "ptr", None)
        if ptr is None:
            counts = torch.bincount(batch)
            ptr = torch.zeros(counts.numel() + 1, dtype=torch.long, device=batch.device)
            ptr[1:] = torch.cumsum(counts, 0)
        topo = Topo.from_sorted_ptr(batch, ptr.to(batch.device))
        data._pool_topo = topo
    return topo
