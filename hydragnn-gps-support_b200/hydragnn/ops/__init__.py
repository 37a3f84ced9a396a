from .topo import Topo, build_csr
from .contract import (
    contract_node,
    contract_edge_w,
    contract_edge_u,
    segment_sum,
    segment_mean,
    gather,
    gather_mul_segment_sum,
    edge_gather_add,
)
from .radius import radius_graph, radius_graph_pbc, MODE_ASE, MODE_CLUSTER
from .attention import attn_varlen, attn_varlen_packed
from .pna import pna_aggregate
from .linear import linear_rows, colsum
from .node import linear_fast, linear_tc, wgrad_rows
