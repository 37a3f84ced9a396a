"""Error of the tensor-core kernels against fp64 as the accumulation gets longer (K of a Linear, rows of a
weight gradient, keys of an attention segment), next to the error of the same op in plain fp32 on the CPU.
Prints |x - fp64|_max / |fp64|_max.      python profiles/precision_probe.py"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "hydragnn-gps-support_b200")]
from hydragnn.ops import attention, node  # noqa: E402
from oracle import kernel_spec as ks  # noqa: E402


def rel(a, truth):
    return float((a.double().cpu() - truth).abs().max() / truth.abs().max())


def main():
    gen = torch.Generator().manual_seed(0)
    print("# linear_tc  y = x w^T, M=1024 N=128:  K  data  gpu  cpu32")
    for K in (32, 128, 256, 384, 1024, 4096):
        for kind in ("randn", "positive"):
            x = torch.randn(1024, K, generator=gen) if kind == "randn" else torch.rand(1024, K, generator=gen)
            w = torch.randn(128, K, generator=gen) if kind == "randn" else torch.rand(128, K, generator=gen)
            truth = x.double() @ w.double().t()
            got = node.linear_tc(x.cuda(), w.cuda())
            print(f"linear K={K:5d} {kind:8s} gpu {rel(got, truth):.2e}  cpu32 {rel(x @ w.t(), truth):.2e}")
    print("# wgrad_rows  dW = a^T b, 128 x 128:  M  data  gpu  cpu32")
    for M in (1024, 8192, 65536, 474078):
        for kind in ("randn", "positive"):
            a = torch.randn(M, 128, generator=gen) if kind == "randn" else torch.rand(M, 128, generator=gen)
            b = torch.randn(M, 128, generator=gen) if kind == "randn" else torch.rand(M, 128, generator=gen)
            truth = a.double().t() @ b.double()
            dw, db = node.wgrad_rows(a.cuda(), b.cuda())
            print(f"wgrad M={M:7d} {kind:8s} gpu {rel(dw, truth):.2e}  cpu32 {rel(a.t() @ b, truth):.2e}  "
                  f"db gpu {rel(db, a.double().sum(0)):.2e} cpu32 {rel(a.sum(0), a.double().sum(0)):.2e}")
    print("# attention H=4 D=32, one segment of L:  L  what  gpu  cpu32")
    H, D = 4, 32
    for L in (128, 1024, 4096, 8192):
        q, k, v, do = (torch.randn(L, H * D, generator=gen) for _ in range(4))
        cu = torch.tensor([0, L], dtype=torch.int32)
        scale = D ** -0.5
        o, lse, *_ = attention._attn_fwd_raw(q.cuda(), k.cuda(), v.cuda(), cu.cuda(), H, scale)
        dq, dk, dv = attention._attn_bwd_raw(do.cuda(), o, lse, q.cuda(), k.cuda(), v.cuda(), cu.cuda(), H, scale)
        o64, lse64 = ks.attn_varlen_fwd(q.double(), k.double(), v.double(), cu, H, scale)
        o32, lse32 = ks.attn_varlen_fwd(q, k, v, cu, H, scale)
        t = ks.attn_varlen_bwd(do.double(), q.double(), k.double(), v.double(), cu, H, scale)
        c = ks.attn_varlen_bwd(do, q, k, v, cu, H, scale)
        print(f"attn L={L:5d} o   gpu {rel(o, o64):.2e} cpu32 {rel(o32, o64):.2e} | lse gpu {rel(lse, lse64):.2e} cpu32 {rel(lse32, lse64):.2e}")
        for name, g, c32, t64 in zip(("dq", "dk", "dv"), (dq, dk, dv), c, t):
            print(f"attn L={L:5d} {name}  gpu {rel(g, t64):.2e} cpu32 {rel(c32, t64):.2e}")


if __name__ == "__main__":
    main()
