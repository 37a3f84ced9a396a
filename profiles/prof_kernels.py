"""Launch the hot kernels at BASELINE config-5 shapes (8 x 1024-atom FeSi crystals: N = 8192 nodes,
E ~ 474k edges grouped by destination, F = 128, 4 heads x 32) a few times, for `ncu` captures:

    python profiles/prof_kernels.py && ncu --set full --clock-control none --import-source on \
        -k regex:'k_contract_node|k_attn' -c 12 -o gpurun_out/prof python profiles/prof_kernels.py
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hydragnn-gps-support_b200"))

from hydragnn import ops  # noqa: E402
from hydragnn.ops import attention, contract  # noqa: E402
from hydragnn.ops.topo import Topo  # noqa: E402


def main(reps=3):
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(0)
    N, deg, F, H, D = 8192, 58, 128, 4, 32
    E = N * deg
    dst = torch.arange(N).repeat_interleave(deg)
    src = (dst // 1024) * 1024 + torch.randint(0, 1024, (E,), generator=gen)
    topo = Topo.from_edge_index(torch.stack([src, dst]).to(dev), N, sorted_side=1)
    x = torch.randn(N, F, generator=gen).to(dev)
    W = torch.randn(E, F, generator=gen).to(dev)
    qkv = torch.randn(N, 3 * H * D, generator=gen).to(dev)
    do = torch.randn(N, H * D, generator=gen).to(dev)
    cu = torch.tensor([0, N], dtype=torch.int32, device=dev)
    topo.csr(0), topo.csr(1)
    c = H * D
    for _ in range(reps):
        contract._node_raw(topo, 1, x, 0, W, None, 1, 1, F)          # CFConv aggregate (streams W)
        contract._node_raw(topo, 0, x, 1, W, None, 1, 1, F)          # its backward w.r.t. x (perm gather)
        contract._edge_w_raw(topo, x, 1, x, 0, None, 1, F)           # its backward w.r.t. W
        o, lse, q, k, v = attention._attn_fwd_raw(qkv[:, :c], qkv[:, c:2 * c], qkv[:, 2 * c:], cu, H, D ** -0.5)
        attention._attn_bwd_raw(do, o, lse, q, k, v, cu, H, D ** -0.5)
    torch.cuda.synchronize()
    print("ok")


if __name__ == "__main__":
    main()
