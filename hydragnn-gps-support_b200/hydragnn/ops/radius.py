"""Batched radius-graph construction on the GPU (hgk_radius_graph_*).

Replaces, for a whole mini-batch at once,
  * RadiusGraphPBC.__call__ = ase.neighborlist.neighbor_list + self-loop removal +
    nearest-first in-degree cap  (reference preprocess/graph_samples_checks_and_updates.py:146-282)
  * PyG RadiusGraph / RadiusInteractionGraph -> torch_cluster.radius_graph
    (same file :109-114, models/SCFStack.py:59-61)
Edges come back grouped by destination and ordered by (distance, source,
shift) inside a destination, which is the order `_limit_neighbors` produces
(np.lexsort((length, dst)), :259) for non-tied lengths.
"""
import torch

from .. import _hgk

MODE_ASE = 0
MODE_CLUSTER = 1


def radius_graph(pos, ptr, r, max_num_neighbors=0, loop=False, cell=None, pbc=None,
                 mode=None, return_rowptr=False):
    """pos [N,3] fp32 cuda; ptr [B+1] node offsets; cell [B,3,3], pbc [B,3] bool or None.

    Returns dict(edge_index int64 [2,E] (row 0 = source, row 1 = destination),
                 shift_int int32 [E,3], dist fp64 [E], rowptr int32 [N+1]).
    One device->host read (the edge count) is unavoidable: outputs are sized by it.
    """
    if not pos.is_cuda:
        raise _hgk.HgkError("radius_graph runs on the GPU only")
    L = _hgk.lib()
    if mode is None:
        mode = MODE_ASE if cell is not None else MODE_CLUSTER
    pos = pos.detach().to(torch.float32).contiguous()
    N = pos.size(0)
    ptr32 = ptr.to(device=pos.device, dtype=torch.int32).contiguous()
    B = ptr32.numel() - 1
    cell_t = pbc_t = None
    if cell is not None:
        cell_t = cell.detach().to(device=pos.device, dtype=torch.float32).reshape(B, 3, 3).contiguous()
        pbc_t = torch.as_tensor(pbc, device=pos.device).reshape(B, 3).to(torch.uint8).contiguous()
    dev = pos.device
    ws_bytes = L.hgk_radius_graph_workspace_bytes(N, B)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    deg_raw = torch.empty(N, dtype=torch.int32, device=dev)
    deg = torch.empty(N, dtype=torch.int32, device=dev)
    _hgk.call("hgk_radius_graph_count", _hgk.ptr(pos), _hgk.ptr(ptr32), _hgk.ptr(cell_t), _hgk.ptr(pbc_t),
              N, B, float(r), int(max_num_neighbors or 0), int(bool(loop)), mode,
              _hgk.ptr(deg_raw), _hgk.ptr(deg), _hgk.ptr(ws), ws_bytes, cost=(N * 56, 0))
    rowptr_raw = torch.zeros(N + 1, dtype=torch.int32, device=dev)
    rowptr = torch.zeros(N + 1, dtype=torch.int32, device=dev)
    rowptr_raw[1:] = torch.cumsum(deg_raw, 0)
    rowptr[1:] = torch.cumsum(deg, 0)
    totals = torch.stack([rowptr_raw[-1], rowptr[-1]]).tolist()  # the one host sync
    E_raw, E = int(totals[0]), int(totals[1])
    src = torch.empty(E, dtype=torch.int32, device=dev)
    dst = torch.empty(E, dtype=torch.int32, device=dev)
    shift = torch.empty((E, 3), dtype=torch.int32, device=dev)
    dist = torch.empty(E, dtype=torch.float64, device=dev)
    cand = torch.empty(max(E_raw, 1) * 24, dtype=torch.uint8, device=dev)
    _hgk.call("hgk_radius_graph_fill", _hgk.ptr(pos), _hgk.ptr(ptr32), _hgk.ptr(cell_t), _hgk.ptr(pbc_t),
              N, B, float(r), int(max_num_neighbors or 0), int(bool(loop)), mode,
              _hgk.ptr(rowptr_raw), _hgk.ptr(rowptr), _hgk.ptr(src), _hgk.ptr(dst),
              _hgk.ptr(shift), _hgk.ptr(dist), _hgk.ptr(cand), cand.numel(), _hgk.ptr(ws), ws_bytes,
              cost=(N * 24 + E * 28 + E_raw * 48, 0))
    out = {
        "edge_index": torch.stack([src, dst]).to(torch.int64),
        "src": src,
        "dst": dst,
        "shift_int": shift,
        "dist": dist,
        "rowptr": rowptr,
    }
    if cell_t is not None:
        # edge_shifts = S(float) @ cell   (reference :233-236)
        g_of_edge = torch.bucketize(dst.long(), ptr32[1:].long(), right=True)
        out["edge_shifts"] = torch.einsum("ek,ekd->ed", shift.to(torch.float32), cell_t[g_of_edge])
    return out
