"""Time hgk_wgrad_tn / hgk_schnet_filter_bwd at config-5 shape."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hydragnn-gps-support_b200"))
from hydragnn.ops import filter as filt  # noqa: E402


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
    E, F, G, R, cutoff = 474078, 128, 50, 8, 5.0
    gen = torch.Generator().manual_seed(0)
    dist = (torch.rand(E, generator=gen) * 4.9 + 0.05).to(dev)
    raw = (torch.rand(E, R, generator=gen) * 0.2).to(dev)
    offset = torch.linspace(0.0, cutoff, G).to(dev)
    coeff = -0.5 / float(offset[1] - offset[0]) ** 2
    a = torch.randn(E, F, generator=gen).to(dev)
    h = torch.randn(E, F, generator=gen).to(dev)
    w1t = (torch.randn(F, F, generator=gen) / 11).to(dev)
    print(f"wgrad [E,128]x[E,128]: {timed(lambda: filt._wgrad_raw(a, dist, cutoff, b=h)):.3f} ms")
    print(f"wgrad [E,128]x gauss : {timed(lambda: filt._wgrad_raw(a, None, cutoff, gauss=(dist, raw, offset, coeff))):.3f} ms")
    print(f"filter_bwd           : {timed(lambda: filt._filter_bwd_raw(a, h, dist, cutoff, w1t)):.3f} ms")


if __name__ == "__main__":
    main()
