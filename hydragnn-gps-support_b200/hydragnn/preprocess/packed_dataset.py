"""A split of the dataset packed into device arenas, with the mini-batch collate on the device.

The reference collates every mini-batch on the host (`Batch.from_data_list`: one torch.cat per field
over the samples of the batch, hydragnn/preprocess/load_data.py:225-326) and then copies it to the GPU.
Here every field of every sample is uploaded ONCE into a packed row matrix (node fields [sum N_i, .],
edge fields [sum E_i, .], ragged targets, per-graph fields) with per-sample row offsets; a mini-batch
is then `collate(indices)`: per field one `hgk_collate_rows` launch that gathers the samples' row
ranges (plus `hgk_collate_edges`, which also shifts the sample-local node ids by the batch's node
offsets, and the `batch` vector written by the node-field gather).  Per batch the host only computes
the output offsets from the stored sample sizes (numpy) and uploads those few integers.

The result equals `Batch.from_data_list([samples[i] for i in indices]).to(device)` field by field,
bit for bit (tests/test_packed_dataset_gpu.py).
"""
import numpy as np
import torch

from .. import _hgk
from ..graph import Batch

_NODE = ("x", "pos", "pe", "forces", "z")
_EDGE = ("edge_attr", "edge_shifts", "rel_pe")
_RAGGED = ("y",)                                   # rows per sample vary (node-level targets)
_GRAPH = ("energy", "cell", "pbc", "y_loc", "dataset_name")


class _Field:
    """Row matrix [R, W] on the device + per-sample row offsets on the host."""

    def __init__(self, parts, device):
        self.tail = tuple(parts[0].shape[1:]) if parts[0].dim() > 1 else ()
        width = 1
        for t in self.tail:
            width *= int(t)
        rows = [p.reshape(p.size(0), width) for p in parts]          # explicit width: samples may have 0 rows
        self.dtype = parts[0].dtype
        self.offsets = np.zeros(len(rows) + 1, dtype=np.int64)
        np.cumsum([r.size(0) for r in rows], out=self.offsets[1:])
        arena = torch.cat(rows, dim=0).contiguous()
        if arena.element_size() * max(arena.size(1), 1) % 4:
            raise ValueError(f"row width of {arena.dtype} x {arena.size(1)} is not a multiple of 4 bytes")
        self.arena = arena.to(device)
        self.row_bytes = arena.element_size() * arena.size(1)


class PackedDataset(torch.utils.data.Dataset):
    """`samples`: Data objects of one split.  `__getitem__` returns the index (so that a DataLoader
    with `collate_fn=self.collate` hands over lists of indices); `collate(indices)` builds the batch."""

    def __init__(self, samples, device):
        self.device = torch.device(device)
        self.samples = samples                     # host copies: config_utils reads sizes / degrees from them
        self.num_samples = len(samples)
        self.fields, self.kind = {}, {}
        keys = set()
        for d in samples:
            keys.update(d.keys())
        for k in sorted(keys):
            vals = [getattr(d, k, None) for d in samples]
            if any(v is None for v in vals) or not torch.is_tensor(vals[0]) or k == "edge_index":
                continue
            if k in _NODE:
                kind = "node"
            elif k in _EDGE:
                kind = "edge"
            elif k in _RAGGED:
                kind = "ragged"
            elif k in _GRAPH:
                kind = "graph"
                if k == "cell":
                    vals = [v.reshape(1, 3, 3) for v in vals]
                elif k == "pbc":
                    vals = [v.reshape(1, 3).to(torch.int32) for v in vals]     # bool rows are 3 bytes: widen
                elif k in ("y_loc", "dataset_name"):
                    vals = [v.reshape(1, -1) for v in vals]
                else:
                    vals = [v.reshape(1, -1) if v.dim() else v.reshape(1, 1) for v in vals]
            else:
                continue
            if kind == "graph" and vals[0].element_size() * vals[0][0].numel() % 4:
                continue
            self.fields[k] = _Field(vals, self.device)
            self.kind[k] = kind
        self.num_nodes = np.array([d.num_nodes for d in samples], dtype=np.int64)
        self.node_off = np.zeros(len(samples) + 1, dtype=np.int64)
        np.cumsum(self.num_nodes, out=self.node_off[1:])
        self.has_edges = all(getattr(d, "edge_index", None) is not None for d in samples) and len(samples) > 0
        if self.has_edges:
            pairs = [d.edge_index.t().contiguous().to(torch.int64) for d in samples]
            self.edge_off = np.zeros(len(samples) + 1, dtype=np.int64)
            np.cumsum([p.size(0) for p in pairs], out=self.edge_off[1:])
            self.edge_pairs = torch.cat(pairs, dim=0).contiguous().to(self.device)

    def __len__(self):
        return self.num_samples

    def __getitem__(self, idx):
        return int(idx)

    # -------------------------------------------------------------------------------------
    def _upload(self, *arrays):
        """Small host integer arrays -> one device int64 tensor (one H2D copy per batch)."""
        flat = np.concatenate([np.asarray(a, dtype=np.int64).ravel() for a in arrays])
        dev = torch.from_numpy(flat).to(self.device, non_blocking=True)
        out, at = [], 0
        for a in arrays:
            n = int(np.asarray(a).size)
            out.append(dev[at:at + n])
            at += n
        return out

    def collate(self, indices):
        idx = np.asarray(list(indices), dtype=np.int64)
        B = int(idx.size)
        out = Batch()
        out.num_graphs = B
        sizes = self.num_nodes[idx]
        node_ptr = np.zeros(B + 1, dtype=np.int64)
        np.cumsum(sizes, out=node_ptr[1:])
        out.graph_sizes = [int(n) for n in sizes]          # host copy, as Batch.from_data_list keeps it
        plan = {"node": (node_ptr, self.node_off[idx])}
        if self.has_edges:
            e_len = self.edge_off[idx + 1] - self.edge_off[idx]
            edge_ptr = np.zeros(B + 1, dtype=np.int64)
            np.cumsum(e_len, out=edge_ptr[1:])
            plan["edge"] = (edge_ptr, self.edge_off[idx])
        for k, f in self.fields.items():
            if self.kind[k] in ("ragged", "graph"):
                lens = f.offsets[idx + 1] - f.offsets[idx]
                ptr = np.zeros(B + 1, dtype=np.int64)
                np.cumsum(lens, out=ptr[1:])
                plan[k] = (ptr, f.offsets[idx])
        names = list(plan)
        dev = self._upload(*[a for n in names for a in plan[n]])
        dplan = {n: (dev[2 * i], dev[2 * i + 1]) for i, n in enumerate(names)}

        out.ptr = dplan["node"][0]
        total_nodes = int(node_ptr[-1])
        out.batch = torch.empty(total_nodes, dtype=torch.int64, device=self.device)
        wrote_batch = False
        for k, f in self.fields.items():
            kind = self.kind[k]
            key = kind if kind in ("node", "edge") else k
            ptr_h = plan[key][0]
            ptr_d, off_d = dplan[key]
            rows = int(ptr_h[-1])
            width = f.row_bytes // f.arena.element_size()
            res = torch.empty((rows, width), dtype=f.dtype, device=self.device)
            seg = None
            if kind == "node" and not wrote_batch:
                seg, wrote_batch = out.batch, True
            _hgk.call("hgk_collate_rows", _hgk.ptr(res), _hgk.ptr(f.arena), _hgk.ptr(ptr_d), _hgk.ptr(off_d), B,
                      f.row_bytes, _hgk.ptr(seg), rows, cost=(2 * rows * f.row_bytes + 16 * B, 0))
            if kind == "graph":
                if k == "cell":
                    res = res.reshape(B, 3, 3)
                elif k == "pbc":
                    res = res.reshape(B, 3).to(torch.bool)
                elif k == "energy":
                    res = res.reshape(-1)
            elif f.tail:
                res = res.reshape((rows,) + f.tail)
            elif kind in ("node", "edge", "ragged") and f.tail == () and width == 1:
                res = res.reshape(rows)
            setattr(out, k, res)
        if not wrote_batch:
            out.batch = torch.repeat_interleave(torch.arange(B, device=self.device), dplan["node"][0][1:] - dplan["node"][0][:-1],
                                                output_size=total_nodes)
        if self.has_edges:
            E = int(plan["edge"][0][-1])
            ei = torch.empty((2, E), dtype=torch.int64, device=self.device)
            _hgk.call("hgk_collate_edges", _hgk.ptr(ei), _hgk.ptr(self.edge_pairs), _hgk.ptr(dplan["edge"][0]),
                      _hgk.ptr(dplan["edge"][1]), _hgk.ptr(dplan["node"][0]), B, E, cost=(32 * E + 24 * B, 0))
            out.edge_index = ei
        return out
