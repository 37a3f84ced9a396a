"""Edge geometry on the GPU path.

Same contract as the reference's get_edge_vectors_and_lengths
(hydragnn/utils/model/operations.py:21-36): vectors = pos[receiver] - pos[sender]
+ shifts, lengths = |vectors|, optional normalisation by (length + eps).  The two
position gathers go through the contraction family so that their backward (and the
backward of that backward, needed for forces) is a deterministic segmented sum
instead of index_select's atomic scatter.
"""
import torch

from ...graph import topo_of
from ...ops import gather
from ...ops import geometry
from ...ops.node import device_ok


def get_edge_vectors_and_lengths(positions, edge_index, shifts, normalize=False, eps=1e-9):
    topo = topo_of(edge_index, positions.size(0))
    if device_ok(positions) and positions.size(-1) == 3 and (shifts is None or shifts.dtype == torch.float32):
        # one operator with hand-written first and second derivatives (ops/geometry.py)
        vec, length, unit = geometry.edge_geometry(positions, topo, shifts, eps, want_unit=normalize)
        return (unit if normalize else vec), length.unsqueeze(-1)
    vectors = gather(positions, topo, 1) - gather(positions, topo, 0)
    if shifts is not None:
        vectors = vectors + shifts
    lengths = torch.linalg.norm(vectors, dim=-1, keepdim=True)
    if normalize:
        return vectors / (lengths + eps), lengths
    return vectors, lengths
