"""Every C-ABI entry point, timed alone at the shapes of the BASELINE configs, against the measured
peaks of MEASURED_PEAKS.json.  CUDA events around each call (the `_hgk.stats` machinery bench.py
uses), an L2 flush (256 MB write) between calls, algorithmic bytes / flops from the same `cost=`
formulas the product passes.  Prints a markdown table.

    python profiles/kernel_rooflines.py [reps] > profiles/r01_kernel_rooflines.md
"""
import json
import math
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hydragnn-gps-support_b200"))

from hydragnn import _hgk  # noqa: E402
from hydragnn.ops import attention, contract, filter as filt, node, pna, radius  # noqa: E402
from hydragnn.ops.topo import Topo, build_csr  # noqa: E402

DEV = torch.device("cuda:0")
ROWS = []


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return p["hbm_gbs"], p["bf16_tflops"], "measured (burst)"
    except Exception:
        return 7700.0, 2250.0, "nominal"


def run(label, shape, fn, reps, flush):
    """Time `fn` `reps` times with a flush before each call; one table row per entry point used."""
    fn()
    torch.cuda.synchronize()
    _hgk.stats.reset()
    _hgk.stats.timing = True
    for _ in range(reps):
        flush.add_(1.0)
        fn()
    torch.cuda.synchronize()
    summ = _hgk.stats.summary()
    _hgk.stats.timing = False
    for name, s in summ.items():
        n = s["calls"]
        ROWS.append(dict(label=label, shape=shape, entry=name, ms=s["ms"] / n, bytes=s["bytes"] / n,
                         flops=s["flops"] / n, dram=s["dram_bytes"] / n))


def fesi_topology(gen, N=8192, deg=58, per_graph=1024):
    E = N * deg
    dst = torch.arange(N).repeat_interleave(deg)
    src = (dst // per_graph) * per_graph + torch.randint(0, per_graph, (E,), generator=gen)
    return Topo.from_edge_index(torch.stack([src, dst]).to(DEV), N, sorted_side=1), E


def main():
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
    gen = torch.Generator().manual_seed(0)
    flush = torch.zeros(64 * 1024 * 1024, device=DEV)            # 256 MB > 126 MB L2
    hbm, tf, src_peak = peaks()

    # ---- config 5 (SchNet+GPS, FeSi-1024 x 8): N = 8192, E = 475k, F = 128, 4 heads x 32 ----------
    N, F, H, D = 8192, 128, 4, 32
    topo, E = fesi_topology(gen)
    topo.csr(0), topo.csr(1)
    shape5 = f"cfg5 N={N} E={E} F={F}"
    x = torch.randn(N, F, generator=gen).to(DEV)
    W = torch.randn(E, F, generator=gen).to(DEV)
    run("CFConv aggregate (W streamed, rows gathered by src)", shape5,
        lambda: contract._node_raw(topo, 1, x, 0, W, None, 1, 1, F), reps, flush)
    run("its d/dx (scatter side: perm gather)", shape5,
        lambda: contract._node_raw(topo, 0, x, 1, W, None, 1, 1, F), reps, flush)
    run("its d/dW", shape5, lambda: contract._edge_w_raw(topo, x, 1, x, 0, None, 1, F), reps, flush)
    run("segment sum of edge rows (no gather)", shape5,
        lambda: contract._node_raw(topo, 1, None, 0, W, None, 1, 1, F), reps, flush)
    key = topo.idx[1]
    run("CSR build (dst-major, already grouped)", shape5, lambda: build_csr(key, N), reps, flush)
    key0 = topo.idx[0]
    run("CSR build (src-major, scattered)", shape5, lambda: build_csr(key0, N), reps, flush)

    d = (torch.rand(E, generator=gen) * 4.9 + 0.05).to(DEV)
    raw = torch.rand(E, 8, generator=gen).to(DEV)
    G = 50
    offset = torch.linspace(0, 5.0, G).to(DEV)
    coeff = -0.5 / float(offset[1] - offset[0]) ** 2
    w0g = (torch.randn(F, G, generator=gen) * 0.1).to(DEV)
    w0r = (torch.randn(F, 8, generator=gen) * 0.1).to(DEV)
    b0 = torch.zeros(F, device=DEV)
    w1 = (torch.randn(F, F, generator=gen) * 0.1).to(DEV)
    b1 = torch.zeros(F, device=DEV)
    GP = (G + 3) // 4 * 4
    w0p = torch.zeros((F, 64), device=DEV)
    w0p[:, :G] = w0g
    w0p[:, GP:GP + 8] = w0r
    out = {}

    def f_fwd():
        out["w"], out["h"] = filt._filter_fwd_raw(d, raw, offset, coeff, 5.0, w0p, b0, w1, b1)
    run("filter network forward (gauss+raw -> Linear -> ssp -> Linear -> envelope)", shape5, f_fwd, reps, flush)
    gw = torch.randn(E, F, generator=gen).to(DEV)
    w1t = w1.t().contiguous()

    def f_bwd():
        out["dz0"] = filt._filter_bwd_raw(gw, out["h"], d, 5.0, w1t)
    run("filter network backward (dZ0)", shape5, f_bwd, reps, flush)
    run("edge-row weight gradient dW1 (A scaled by envelope, B = h)", shape5,
        lambda: filt._wgrad_raw(gw, d, 5.0, b=out["h"]), reps, flush)
    run("edge-row weight gradient dW0 (B = gaussians | raw rebuilt on the fly)", shape5,
        lambda: filt._wgrad_raw(out["dz0"], None, 5.0, gauss=(d, raw, offset, coeff)), reps, flush)

    qkv = torch.randn(N, 3 * H * D, generator=gen).to(DEV)
    do = torch.randn(N, H * D, generator=gen).to(DEV)
    c = H * D
    for name, cu in [("whole batch (1 x 8192)", torch.tensor([0, N], dtype=torch.int32, device=DEV)),
                     ("per graph (8 x 1024)", torch.arange(0, N + 1, 1024, dtype=torch.int32, device=DEV))]:
        st = {}

        def a_fwd():
            st["r"] = attention._attn_fwd_raw(qkv[:, :c], qkv[:, c:2 * c], qkv[:, 2 * c:], cu, H, D ** -0.5)
        run(f"GPS attention forward, {name}", f"cfg5 N={N} H={H} D={D}", a_fwd, reps, flush)
        o, lse, q, k, v = st["r"]
        run(f"GPS attention backward, {name}", f"cfg5 N={N} H={H} D={D}",
            lambda: attention._attn_bwd_raw(do, o, lse, q, k, v, cu, H, D ** -0.5), reps, flush)

    # node-row kernels of the GPS block at config 5 (N = 8192 rows, C = 128)
    shape_n = f"cfg5 N={N} C={F}"
    w_qkv = (torch.randn(3 * F, F, generator=gen) * 0.1).to(DEV)
    w_sq = (torch.randn(F, F, generator=gen) * 0.1).to(DEV)
    w_ff2 = (torch.randn(F, 2 * F, generator=gen) * 0.1).to(DEV)
    b_qkv = torch.zeros(3 * F, device=DEV)
    u2 = torch.randn(N, 2 * F, generator=gen).to(DEV)
    mask = (torch.rand(N, F, generator=gen) < 0.75).to(torch.uint8).to(DEV)
    run("Linear 128 -> 128 (+ bias)", shape_n, lambda: node.linear_tc(x, w_sq, b1), reps, flush)
    run("QKV projection 128 -> 384 (+ bias)", shape_n, lambda: node.linear_tc(x, w_qkv, b_qkv), reps, flush)
    run("FFN second Linear 256 -> 128 (+ bias, dropout mask, residual)", shape_n,
        lambda: node.linear_tc(u2, w_ff2, b1, mask=mask, scale=4.0 / 3.0, residual=x), reps, flush)
    run("input gradient of the QKV projection (transposed weight, K = 384, + residual)", shape_n,
        lambda: node.linear_tc(qkv, w_qkv, trans_w=True, residual=x), reps, flush)
    run("weight + bias gradient of the QKV projection ([384 x 128] over 8192 rows)", shape_n,
        lambda: node.wgrad_rows(qkv, x), reps, flush)
    st_bn = {}

    def bn_s():
        st_bn["p"] = node.BnParams(*node.bn_stats(x), b1 + 1.0, b1)
    run("BatchNorm batch statistics", shape_n, bn_s, reps, flush)
    run("BatchNorm apply (sum of two normalised inputs)", shape_n,
        lambda: node.bn_apply(x, st_bn["p"], do, st_bn["p"]), reps, flush)
    run("BatchNorm backward (two inputs, one dropout mask)", shape_n,
        lambda: node.bn_bwd(do, x, st_bn["p"], do, st_bn["p"], mask_b=mask, scale_b=4.0 / 3.0), reps, flush)

    # periodic radius graph of config 5: 8 jittered BCC 8x8x8 cells, r = 5.0
    a = 2.8
    base = torch.tensor([[i, j, k] for i in range(8) for j in range(8) for k in range(8)], dtype=torch.float32)
    cellpos = torch.cat([base, base + 0.5]) * a
    pos = torch.cat([cellpos + torch.randn(1024, 3, generator=gen) * 0.05 for _ in range(8)]).to(DEV)
    ptr = torch.arange(0, 8193, 1024)
    cell = (torch.eye(3) * 8 * a).repeat(8, 1, 1)
    pbcf = torch.ones(8, 3, dtype=torch.bool)
    run("periodic radius graph (ASE semantics, cap 64)", "cfg5 8 x 1024 atoms, r=5.0 -> 475k edges",
        lambda: radius.radius_graph(pos, ptr, 5.0, 64, False, cell=cell, pbc=pbcf), max(reps // 2, 2), flush)

    # ---- config 4 (PAINN forces, OC20-shaped): N = 2560, E = 1.1e5, three vector channels -----------
    N4, deg4 = 2560, 43
    E4 = N4 * deg4
    dst4 = torch.arange(N4).repeat_interleave(deg4)
    src4 = (dst4 // 80) * 80 + torch.randint(0, 80, (E4,), generator=gen)
    topo4 = Topo.from_edge_index(torch.stack([src4, dst4]).to(DEV), N4, sorted_side=1)
    topo4.csr(0), topo4.csr(1)
    shape4 = f"cfg4 N={N4} E={E4} F=128 C=3"
    v3 = torch.randn(N4, 3, F, generator=gen).to(DEV)
    s1 = torch.randn(N4, F, generator=gen).to(DEV)
    W4 = torch.randn(E4, F, generator=gen).to(DEV)
    U4 = torch.randn(E4, 3, generator=gen).to(DEV)
    run("PaiNN vector message: sum_e v[j,c,:] * fw[e,:]", shape4,
        lambda: contract._node_raw(topo4, 1, v3, 0, W4, None, 3, 3, F), reps, flush)
    run("PaiNN direction message: sum_e phi[j,:] * fw[e,:] * u[e,c]", shape4,
        lambda: contract._node_raw(topo4, 1, s1, 0, W4, U4, 3, 3, F), reps, flush)
    run("its d/dfw (edge_w, 3 channels)", shape4,
        lambda: contract._edge_w_raw(topo4, v3, 1, s1, 0, U4, 3, F), reps, flush)
    run("its d/du (edge_u: dot over F per channel)", shape4,
        lambda: contract._edge_u_raw(topo4, v3, 1, s1, 0, W4, 3, F), reps, flush)

    # ---- config 3 (EGNN, ZINC-shaped) edge-MLP input and a PNA layer at a width that fills the GPU -----
    N3, deg3 = 1485 * 16, 2
    E3 = N3 * deg3
    dst3 = torch.arange(N3).repeat_interleave(deg3)
    src3 = torch.randint(0, N3, (E3,), generator=gen)
    topo3 = Topo.from_edge_index(torch.stack([src3, dst3]).to(DEV), N3, sorted_side=1)
    topo3.csr(0), topo3.csr(1)
    P3 = torch.randn(N3, F, generator=gen).to(DEV)
    Q3 = torch.randn(N3, F, generator=gen).to(DEV)
    run("EGNN edge-MLP input x[row]Wa + x[col]Wb (split-weight form)", f"cfg3 x16: N={N3} E={E3} F=128",
        lambda: contract._edge_gather_add_raw(topo3, P3, 0, Q3, 1), reps, flush)
    Np, degp = 32768, 16
    Ep = Np * degp
    dstp = torch.arange(Np).repeat_interleave(degp)
    srcp = torch.randint(0, Np, (Ep,), generator=gen)
    topop = Topo.from_edge_index(torch.stack([srcp, dstp]).to(DEV), Np, sorted_side=1)
    topop.csr(0), topop.csr(1)
    Pp = torch.randn(Np, 64, generator=gen).to(DEV)
    Qp = torch.randn(Np, 64, generator=gen).to(DEV)
    stp = {}

    def p_fwd():
        stp["r"] = pna._pna_fwd_raw(Pp, Qp, None, topop)
    run("PNA aggregate forward (mean|min|max|std in one pass)", f"N={Np} E={Ep} F=64", p_fwd, reps, flush)
    outp, cmin, cmax = stp["r"]
    gout = torch.randn(Np, 256, generator=gen).to(DEV)
    run("PNA aggregate backward", f"N={Np} E={Ep} F=64",
        lambda: pna._pna_bwd_raw(gout, outp, cmin, cmax, Pp, Qp, None, topop), reps, flush)

    print(f"Peaks: HBM {hbm:.0f} GB/s, dense bf16 {tf:.0f} TFLOP/s ({src_peak}); 3xTF32 ceiling = bf16 / 6.")
    print(f"{reps} timed calls each, 256 MB L2 flush before every call, CUDA events on the launch stream.\n")
    print("| what | shape | entry point | ms | alg. GB/s | compulsory GB/s | frac HBM | TFLOP/s | frac of 3xTF32 ceiling |")
    print("|---|---|---|---|---|---|---|---|---|")
    for r in ROWS:
        gbs = r["bytes"] / r["ms"] / 1e6 if r["ms"] else 0.0
        cgbs = r["dram"] / r["ms"] / 1e6 if r["ms"] else 0.0
        tfl = r["flops"] / r["ms"] / 1e9 if r["ms"] else 0.0
        print(f"| {r['label']} | {r['shape']} | `{r['entry']}` | {r['ms']:.4f} | {gbs:.0f} | {cgbs:.0f} | "
              f"{cgbs / hbm:.2f} | {tfl:.1f} | {(tfl / (tf / 6)):.2f} |" if r["flops"] else
              f"| {r['label']} | {r['shape']} | `{r['entry']}` | {r['ms']:.4f} | {gbs:.0f} | {cgbs:.0f} | "
              f"{cgbs / hbm:.2f} | – | – |")


if __name__ == "__main__":
    main()
