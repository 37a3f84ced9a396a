"""Where the cycles of the pipelined dK/dV attention kernel go: clock64 deltas accumulated by thread 0
of CTA (0,0) in a -DHGK_KV_PROFILE build of the same sources (built into profiles/_build/, loaded
through HGK_LIB).  Slots: 0 wait for the S^T/dP^T MMAs, 1 TMEM loads, 2 exp2 / split math,
3 TMEM stores, 4 staging the tiles of the group's next step, 5 proxy fence + barrier arrive,
6 issuing the global loads of the step after next, 7 loop overhead.

    python profiles/kv_phase_profile.py build      # here (nvcc cross-compiles)
    python profiles/kv_phase_profile.py            # on the GPU box
"""
import ctypes
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "hydragnn-gps-support_b200", "csrc")
OUT = os.path.join(ROOT, "profiles", "_build", "libhgk_prof.so")


def build():
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    srcs = sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))
    cmd = ["nvcc", "-O3", "-std=c++17", "-lineinfo", "-gencode", "arch=compute_100a,code=sm_100a", "--shared",
           "-Xcompiler", "-fPIC", "--threads", "0", "-DHGK_KV_PROFILE", "-o", OUT] + srcs
    subprocess.run(cmd, check=True)
    print(OUT)


def main():
    os.environ["HGK_LIB"] = OUT
    sys.path.insert(0, os.path.join(ROOT, "hydragnn-gps-support_b200"))
    import torch
    from hydragnn import _hgk
    from hydragnn.ops import attention

    dev = torch.device("cuda:0")
    N, H, D = 8192, 4, 32
    cu = torch.tensor([0, N], dtype=torch.int32, device=dev)
    gen = torch.Generator().manual_seed(0)
    qkv = torch.randn(N, 3 * H * D, generator=gen).to(dev)
    do = torch.randn(N, H * D, generator=gen).to(dev)
    c = H * D
    q, k, v = qkv[:, :c], qkv[:, c:2 * c], qkv[:, 2 * c:]
    o, lse, *_ = attention._attn_fwd_raw(q, k, v, cu, H, D ** -0.5)
    L = _hgk.lib()
    fn = L.hgk_debug_kv_profile
    fn.argtypes = [ctypes.c_void_p, ctypes.c_int]
    buf = (ctypes.c_ulonglong * 8)()
    attention._attn_bwd_raw(do, o, lse, q, k, v, cu, H, D ** -0.5)
    fn(None, 1)
    reps = 5
    for _ in range(reps):
        attention._attn_bwd_raw(do, o, lse, q, k, v, cu, H, D ** -0.5)
    fn(buf, 0)
    steps = reps * (N // 64)          # thread 0 is in group 0: every other 32-query step
    names = ["wait MMA", "tmem ld", "math", "tmem st", "stage", "fence+arrive", "load issue", "loop"]
    total = sum(buf)
    print(f"cycles per own step of group 0 (= two 32-query steps of wall time; CTA 0,0; {steps} steps): total {total / steps:.0f}")
    for n, c_ in zip(names, buf):
        print(f"  {n:12s} {c_ / steps:8.1f}")


if __name__ == "__main__":
    build() if len(sys.argv) > 1 and sys.argv[1] == "build" else main()
