"""Time the fused SchNet filter kernel at config-5 shape (E = 474k, F = 128, G = 50, R = 8)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hydragnn-gps-support_b200"))
from hydragnn.ops import filter as filt  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    E, F, G, R, cutoff = 474078, 128, 50, 8, 5.0
    gen = torch.Generator().manual_seed(0)
    dist = (torch.rand(E, generator=gen) * 4.9 + 0.05).to(dev)
    raw = (torch.rand(E, R, generator=gen) * 0.2).to(dev)
    offset = torch.linspace(0.0, cutoff, G).to(dev)
    coeff = -0.5 / float(offset[1] - offset[0]) ** 2
    w0p = (torch.randn(F, 64, generator=gen) / 8).to(dev)
    w1 = (torch.randn(F, F, generator=gen) / 11).to(dev)
    b = torch.zeros(F, device=dev)
    fn = lambda: filt._filter_fwd_raw(dist, raw, offset, coeff, cutoff, w0p, b, w1, b)
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    flops = 2 * E * F * (G + R + F)
    print(f"schnet_filter_fwd {ms:.3f} ms  {flops / ms / 1e9:.1f} TFLOP/s alg  {E * (8 * F + 36) / ms / 1e6:.0f} GB/s")


if __name__ == "__main__":
    main()
