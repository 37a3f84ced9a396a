"""Where the cycles of the fused filter-network forward kernel go (thread 0 of CTA 0, clock64, the
-DHGK_KV_PROFILE build made by `python profiles/kv_phase_profile.py build`).  Slots: 0 build X
(gaussians) + TMEM store, 1 barrier, 2 issue MMA0, 3 wait MMA0, 4 ssp epilogue + h store + TMEM
store, 5 barrier + issue MMA1, 6 wait MMA1, 7 output epilogue (w store) + loop."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ["HGK_LIB"] = os.path.join(ROOT, "profiles", "_build", "libhgk_prof.so")
sys.path.insert(0, os.path.join(ROOT, "hydragnn-gps-support_b200"))
import torch  # noqa: E402
from hydragnn import _hgk  # noqa: E402
from hydragnn.ops import filter as filt  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    E, F, G = 475136, 128, 50
    gen = torch.Generator().manual_seed(0)
    d = (torch.rand(E, generator=gen) * 4.9 + 0.05).to(dev)
    raw = torch.rand(E, 8, generator=gen).to(dev)
    offset = torch.linspace(0, 5.0, G).to(dev)
    coeff = -0.5 / float(offset[1] - offset[0]) ** 2
    w0p = (torch.randn(F, 64, generator=gen) * 0.1).to(dev)
    b0 = torch.zeros(F, device=dev)
    w1 = (torch.randn(F, F, generator=gen) * 0.1).to(dev)
    L = _hgk.lib()
    fn = L.hgk_debug_filter_profile
    fn.argtypes = [ctypes.c_void_p, ctypes.c_int]
    buf = (ctypes.c_ulonglong * 8)()
    filt._filter_fwd_raw(d, raw, offset, coeff, 5.0, w0p, b0, w1, b0)
    fn(None, 1)
    reps = 5
    for _ in range(reps):
        filt._filter_fwd_raw(d, raw, offset, coeff, 5.0, w0p, b0, w1, b0)
    fn(buf, 0)
    tiles = reps * ((E // 128 + 147) // 148)
    names = ["build X", "barrier", "issue MMA0", "wait MMA0", "ssp + h store", "barrier + issue MMA1", "wait MMA1",
             "w store + loop"]
    print(f"cycles per 128-edge tile (CTA 0; {tiles} tiles): total {sum(buf) / tiles:.0f}")
    for n, c_ in zip(names, buf):
        print(f"  {n:22s} {c_ / tiles:8.1f}")


if __name__ == "__main__":
    main()
