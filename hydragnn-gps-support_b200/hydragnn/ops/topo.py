"""Edge topology of one mini-batch: endpoint arrays + the CSRs the kernels need.

The reference passes `edge_index` (int64 [2,E]) to torch scatter_add_/index_add_
every time it aggregates (EGCLStack.py:292-298, PAINNStack.py:254-257,
SCFStack.py:285).  Here the grouping work is done once per batch: a `Topo`
holds int32 endpoint arrays for "side 0" (edge_index[0]) and "side 1"
(edge_index[1]) and, per side, a stable CSR (rowptr, perm) built by
hgk_csr_build.  perm is None when the edges are already grouped by that side
(radius graphs come out grouped by destination), which lets the kernels stream
edge rows instead of gathering them.
"""
import torch

from .. import _hgk


def build_csr(key, num_rows):
    """Stable CSR of edge ids grouped by `key` (int32 [E]) -> (rowptr, perm)."""
    E = key.numel()
    L = _hgk.lib()
    rowptr = torch.empty(num_rows + 1, dtype=torch.int32, device=key.device)
    perm = torch.empty(max(E, 1), dtype=torch.int32, device=key.device)
    ws_bytes = L.hgk_csr_build_workspace_bytes(E, num_rows)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=key.device)
    # key read twice, slot + tmp written and read, perm written; counters + rowptr per row
    _hgk.call("hgk_csr_build", _hgk.ptr(key), E, num_rows, _hgk.ptr(rowptr), _hgk.ptr(perm),
              _hgk.ptr(ws), ws_bytes, cost=(E * 28 + num_rows * 12, 0))
    return rowptr, perm[:E]


class Topo:
    """Endpoint arrays and per-side CSR of E edges over `num_nodes` rows.

    sides: 0 = edge_index[0] (the reference's `row` / source / `i`),
           1 = edge_index[1] (`col` / target / `j`).
    A one-sided Topo (pooling: "edges" are nodes, side 0 is the graph id) is
    built with `Topo.from_sorted_ptr`.
    """

    def __init__(self, idx0, idx1, num_nodes, sorted_side=None, num_nodes1=None):
        self.E = int(idx0.numel())
        self.idx = [idx0, idx1]
        self.num_rows = [int(num_nodes), int(num_nodes if num_nodes1 is None else num_nodes1)]
        self._csr = [None, None]
        self.sorted_side = sorted_side
        self.device = idx0.device

    @classmethod
    def from_edge_index(cls, edge_index, num_nodes, sorted_side=None):
        ei = edge_index.to(torch.int32)
        return cls(ei[0].contiguous(), ei[1].contiguous(), num_nodes, sorted_side)

    @classmethod
    def from_sorted_ptr(cls, seg_id, ptr):
        """One-sided topology for sorted segments (node -> graph pooling)."""
        t = cls(seg_id.to(torch.int32).contiguous(), None, ptr.numel() - 1, sorted_side=0)
        t._csr[0] = (ptr.to(torch.int32).contiguous(), None)
        return t

    def csr(self, side):
        if self._csr[side] is None:
            key = self.idx[side]
            if self.sorted_side == side:
                # already grouped (ascending): rowptr[n] = first position with key >= n, identity
                # permutation.  A binary search per row: no device -> host sync (bincount has one)
                bounds = torch.arange(self.num_rows[side] + 1, dtype=key.dtype, device=key.device)
                rowptr = torch.searchsorted(key, bounds, out_int32=True)
                self._csr[side] = (rowptr, None)
            else:
                self._csr[side] = build_csr(key, self.num_rows[side])
        return self._csr[side]

    def set_csr(self, side, rowptr, perm=None):
        self._csr[side] = (rowptr, perm)

    def degree(self, side):
        rowptr, _ = self.csr(side)
        return (rowptr[1:] - rowptr[:-1])
