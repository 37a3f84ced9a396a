"""Time the varlen attention kernels alone at config-5 shape (N = 8192 whole-batch, or 8 x 1024 per
graph), CUDA events, 3 warm-up + 10 timed launches.   python profiles/time_attn.py [graph]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hydragnn-gps-support_b200"))
from hydragnn.ops import attention  # noqa: E402


def timed(fn, reps=10):
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
    N, H, D = 8192, 4, 32
    per_graph = len(sys.argv) > 1 and sys.argv[1] == "graph"
    cu = torch.arange(0, N + 1, 1024, dtype=torch.int32, device=dev) if per_graph else \
        torch.tensor([0, N], dtype=torch.int32, device=dev)
    gen = torch.Generator().manual_seed(0)
    qkv = torch.randn(N, 3 * H * D, generator=gen).to(dev)
    do = torch.randn(N, H * D, generator=gen).to(dev)
    c = H * D
    q, k, v = qkv[:, :c], qkv[:, c:2 * c], qkv[:, 2 * c:]
    scale = D ** -0.5
    o, lse, *_ = attention._attn_fwd_raw(q, k, v, cu, H, scale)
    lens = (cu[1:] - cu[:-1]).double()
    flops = float(4 * c * (lens * lens).sum())
    t_f = timed(lambda: attention._attn_fwd_raw(q, k, v, cu, H, scale))
    t_b = timed(lambda: attention._attn_bwd_raw(do, o, lse, q, k, v, cu, H, scale))
    print(f"segments={'graph' if per_graph else 'batch'} fwd {t_f:.3f} ms ({flops / t_f / 1e9:.1f} TFLOP/s alg) "
          f"bwd {t_b:.3f} ms ({2.5 * flops / t_b / 1e9:.1f} TFLOP/s alg)")


if __name__ == "__main__":
    main()
