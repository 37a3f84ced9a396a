"""Launch the hot kernels at BASELINE config-5 shapes (8 x 1024-atom FeSi crystals: N = 8192 nodes,
E ~ 474k edges grouped by destination, F = 128, 4 heads x 32) a few times, for `ncu` captures:

    python profiles/prof_kernels.py && ncu --set full --clock-control none --import-source on \
        -k regex:'k_contract_node|k_attn|k_schnet|k_wgrad' -c 14 -o gpurun_out/prof python profiles/prof_kernels.py
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hydragnn-gps-support_b200"))

from hydragnn import ops  # noqa: E402
from hydragnn.ops import attention, contract, node, filter as filt  # noqa: E402
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
    G = 50
    d = (torch.rand(E, generator=gen) * 4.9 + 0.05).to(dev)
    raw = torch.rand(E, 8, generator=gen).to(dev)
    offset = torch.linspace(0, 5.0, G).to(dev)
    coeff = -0.5 / float(offset[1] - offset[0]) ** 2
    w0p = torch.zeros(F, 64, device=dev)
    w0p[:, :G] = torch.randn(F, G, generator=gen).to(dev) * 0.1
    w0p[:, 52:60] = torch.randn(F, 8, generator=gen).to(dev) * 0.1
    b0 = torch.zeros(F, device=dev)
    w1 = (torch.randn(F, F, generator=gen) * 0.1).to(dev)
    for _ in range(reps):
        fw, fh = filt._filter_fwd_raw(d, raw, offset, coeff, 5.0, w0p, b0, w1, b0)      # filter network forward
        dz0 = filt._filter_bwd_raw(W, fh, d, 5.0, w1.t().contiguous())                  # its backward (dZ0)
        filt._wgrad_raw(W, d, 5.0, b=fh)                                                # dW1 (envelope-scaled A)
        filt._wgrad_raw(dz0, None, 5.0, gauss=(d, raw, offset, coeff))                  # dW0 (gaussians rebuilt)
        contract._node_raw(topo, 1, x, 0, W, None, 1, 1, F)          # CFConv aggregate (streams W)
        contract._node_raw(topo, 0, x, 1, W, None, 1, 1, F)          # its backward w.r.t. x (perm gather)
        contract._edge_w_raw(topo, x, 1, x, 0, None, 1, F)           # its backward w.r.t. W
        o, lse, q, k, v = attention._attn_fwd_raw(qkv[:, :c], qkv[:, c:2 * c], qkv[:, 2 * c:], cu, H, D ** -0.5)
        attention._attn_bwd_raw(do, o, lse, q, k, v, cu, H, D ** -0.5)
        # node-row kernels of the GPS block: QKV projection, its input gradient, its weight gradient, BatchNorm
        win = (torch.randn(3 * c, c, generator=gen) * 0.1).to(dev)
        y = node.linear_tc(x, win, torch.zeros(3 * c, device=dev))
        node.linear_tc(y, win, trans_w=True, residual=x)
        node.wgrad_rows(y, x)
        mean, rstd = node.bn_stats(x)
        node.bn_apply(x, node.BnParams(mean, rstd, None, None))
    torch.cuda.synchronize()
    print("ok")


if __name__ == "__main__":
    main()
