"""Time the contraction kernels at config-5 shape (N = 8192, E = 58 N, F = 128)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hydragnn-gps-support_b200"))
from hydragnn.ops import contract  # noqa: E402
from hydragnn.ops.topo import Topo  # noqa: E402


def timed(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(0)
    N, deg, F = 8192, 58, 128
    E = N * deg
    dst = torch.arange(N).repeat_interleave(deg)
    src = (dst // 1024) * 1024 + torch.randint(0, 1024, (E,), generator=gen)
    topo = Topo.from_edge_index(torch.stack([src, dst]).to(dev), N, sorted_side=1)
    topo.csr(0), topo.csr(1)
    x = torch.randn(N, F, generator=gen).to(dev)
    W = torch.randn(E, F, generator=gen).to(dev)
    flush = torch.empty(64 * 1024 * 1024, device=dev)
    alg = E * (8 * F + 4) + N * (4 * F + 4)
    for name, fn in [("aggregate  (dst CSR, streams W)", lambda: contract._node_raw(topo, 1, x, 0, W, None, 1, 1, F)),
                     ("grad wrt x (src CSR, perm)     ", lambda: contract._node_raw(topo, 0, x, 1, W, None, 1, 1, F)),
                     ("segment_sum (W only)           ", lambda: contract._node_raw(topo, 1, None, 0, W, None, 1, 1, F)),
                     ("edge_w     (gather x gather)   ", lambda: contract._edge_w_raw(topo, x, 1, x, 0, None, 1, F))]:
        ms = timed(lambda: (flush.zero_(), fn())[1]) - timed(lambda: flush.zero_())
        print(f"{name}: {ms:.4f} ms  {alg / ms / 1e6:.0f} GB/s algorithmic")


if __name__ == "__main__":
    main()
