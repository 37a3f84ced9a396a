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


def _graphs_with_isolated_nodes(rowptr, graph_of_node, num_graphs):
    isolated = rowptr[1:] == rowptr[:-1]
    count = torch.zeros(num_graphs, dtype=torch.int64, device=rowptr.device)
    count.index_add_(0, graph_of_node, isolated.to(torch.int64))     # integer adds: order independent
    return count > 0, isolated


def radius_graph_pbc(pos, ptr, r, max_num_neighbors=0, loop=False, cell=None, pbc=None,
                     cutoff_multiplier=1.25, max_attempts=3):
    """RadiusGraphPBC.__call__ for a whole batch of samples, including the part of it that is
    per-sample control flow in the reference (graph_samples_checks_and_updates.py:163-221,284-307):
    a sample in which some node receives no edge is searched again with the cutoff x1.25 (at most
    `max_attempts` searches), and nodes still unconnected after the last search get one artificial
    incoming edge from a random other node of the sample (zero shift, appended after the sample's
    regular edges; the source is drawn with np.random.choice in ascending node order, like the
    reference, so seeding numpy reproduces it).

    Only the samples that need it are searched again -- as a sub-batch through the same kernels.
    Returns the dict of `radius_graph` plus `edge_ptr` int64 [B+1] (edge offsets per sample, host)
    and `cutoff` fp64 [B] (the cutoff each sample ended with); `rowptr` is None when artificial
    edges were appended (they are not grouped by destination)."""
    import numpy as np

    if cell is None:
        raise _hgk.HgkError("radius_graph_pbc needs a cell per sample")
    dev = pos.device
    ptr_d = ptr.to(device=dev, dtype=torch.int64)
    B, N = ptr_d.numel() - 1, pos.size(0)
    sizes = ptr_d[1:] - ptr_d[:-1]
    graph_of_node = torch.repeat_interleave(torch.arange(B, device=dev), sizes)
    cell_d = cell.detach().to(device=dev, dtype=torch.float32).reshape(B, 3, 3)
    pbc_d = torch.as_tensor(pbc, device=dev).reshape(B, 3)
    cutoffs = torch.full((B,), float(r), dtype=torch.float64)

    g = radius_graph(pos, ptr_d, r, max_num_neighbors, loop, cell_d, pbc_d, mode=MODE_ASE)
    bad, _ = _graphs_with_isolated_nodes(g["rowptr"], graph_of_node, B)
    if max_attempts <= 1 or not bool(bad.any()):
        g["edge_ptr"] = g["rowptr"].long()[ptr_d].cpu()
        g["cutoff"] = cutoffs
        return g

    # edges of the samples that are complete, by search level; a sample is in exactly one level
    edge_graph = graph_of_node[g["dst"].long()]
    keep = ~bad[edge_graph]
    parts = [dict(src=g["src"][keep].long(), dst=g["dst"][keep].long(), shift=g["shift_int"][keep],
                  dist=g["dist"][keep], art=torch.zeros(int(keep.sum()), dtype=torch.int64, device=dev))]
    todo = torch.nonzero(bad).flatten()                      # global sample ids still incomplete
    cutoff = float(r)
    for attempt in range(1, max_attempts):
        cutoff *= cutoff_multiplier
        cutoffs[todo.cpu()] = cutoff
        sel = torch.zeros(B, dtype=torch.bool, device=dev)
        sel[todo] = True
        nodes = torch.nonzero(sel[graph_of_node]).flatten()  # global node ids of the sub-batch, ascending
        sub_sizes = sizes[todo]
        sub_ptr = torch.zeros(todo.numel() + 1, dtype=torch.int64, device=dev)
        sub_ptr[1:] = torch.cumsum(sub_sizes, 0)
        sub_gon = torch.repeat_interleave(torch.arange(todo.numel(), device=dev), sub_sizes)
        sg = radius_graph(pos[nodes], sub_ptr, cutoff, max_num_neighbors, loop, cell_d[todo], pbc_d[todo],
                          mode=MODE_ASE)
        sub_bad, isolated = _graphs_with_isolated_nodes(sg["rowptr"], sub_gon, todo.numel())
        last = attempt == max_attempts - 1
        keep = torch.ones_like(sg["dst"], dtype=torch.bool) if last else ~sub_bad[sub_gon[sg["dst"].long()]]
        parts.append(dict(src=nodes[sg["src"][keep].long()], dst=nodes[sg["dst"][keep].long()],
                          shift=sg["shift_int"][keep], dist=sg["dist"][keep],
                          art=torch.zeros(int(keep.sum()), dtype=torch.int64, device=dev)))
        if last:
            missing = nodes[isolated].cpu().numpy()          # ascending: sample order, node order inside
            if missing.size:
                ptr_h = ptr_d.cpu().numpy()
                gon_h = graph_of_node.cpu().numpy()
                a_src = []
                for m in missing:
                    lo, hi = int(ptr_h[gon_h[m]]), int(ptr_h[gon_h[m] + 1])
                    local = np.arange(hi - lo)
                    a_src.append(lo + (int(np.random.choice(local[local != m - lo])) if hi - lo > 1 else 0))
                k = len(a_src)
                parts.append(dict(src=torch.tensor(a_src, dtype=torch.int64, device=dev),
                                  dst=torch.from_numpy(missing).to(dev),
                                  shift=torch.zeros((k, 3), dtype=torch.int32, device=dev),
                                  dist=torch.full((k,), cutoff - 1e-8, dtype=torch.float64, device=dev),
                                  art=torch.ones(k, dtype=torch.int64, device=dev)))
        else:
            todo = todo[sub_bad]
            if todo.numel() == 0:
                break

    cat = {k: torch.cat([p[k] for p in parts]) for k in parts[0]}
    edge_graph = graph_of_node[cat["dst"]]
    order = torch.sort(edge_graph * 2 + cat["art"], stable=True).indices
    src, dst = cat["src"][order], cat["dst"][order]
    shift, edge_graph = cat["shift"][order], edge_graph[order]
    has_art = bool(cat["art"].any())
    edge_ptr = torch.zeros(B + 1, dtype=torch.int64, device=dev)
    edge_ptr[1:] = torch.cumsum(torch.bincount(edge_graph, minlength=B), 0)
    rowptr = None
    if not has_art:
        rowptr = torch.zeros(N + 1, dtype=torch.int32, device=dev)
        rowptr[1:] = torch.cumsum(torch.bincount(dst, minlength=N), 0)
    return {
        "edge_index": torch.stack([src, dst]),
        "src": src.to(torch.int32),
        "dst": dst.to(torch.int32),
        "shift_int": shift,
        "dist": cat["dist"][order],
        "rowptr": rowptr,
        "edge_shifts": torch.einsum("ek,ekd->ed", shift.to(torch.float32), cell_d[edge_graph]),
        "edge_ptr": edge_ptr.cpu(),
        "cutoff": cutoffs,
    }
