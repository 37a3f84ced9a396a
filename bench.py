#!/usr/bin/env python
"""Benchmark of the message-passing training step (BASELINE.json: train graphs/sec at 1/2/4/8 B200,
message-pass HBM GB/s vs peak).

    python bench.py --gpus N --steps K --warmup W [--workload NAME] [--impl reference]

One "step" = forward + loss + backward (+ double backward for forces) + gradient all-reduce +
AdamW step on one synthetic mini-batch per GPU.  Default workload: BASELINE config 5 (SchNet + GPS
global attention on 1024-atom periodic FeSi crystals, 8 graphs per GPU, weak scaling), the config
the multi-GPU metric is quoted on; `--workload` selects configs 2-4 (see hydragnn/utils/synthetic.py).
Prints ONE JSON line (rank 0).  `--impl reference` times the CPU restatement of the reference
(oracle/) on the host cores instead -- the only place besides `cpu_baseline` where oracle/ runs.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "hydragnn-gps-support_b200")
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

L2_BYTES = 126 * 1024 * 1024


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default="schnet_gps_fesi1024")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--gps-segments", default="batch", choices=["batch", "graph"],
                    help="batch = the reference's whole-mini-batch attention (SURVEY F3); graph = per-graph varlen")
    ap.add_argument("--graphs-per-gpu", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-kernel-timing", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    return ap.parse_args()


# --------------------------------------------------------------------------- clocks ---
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""

    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.FIELDS}",
                 "--format=csv,noheader,nounits", "-lms", "200"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *exc):
        if self.proc is not None:
            time.sleep(0.25)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for name, val in zip(names, r[2:6]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm)}


# ----------------------------------------------------------------------------- data ---
def build_batch_gpu(name, graphs, device):
    """Collate + (for radius workloads) neighbour lists from the cell-list kernel -> CPU tensors."""
    from hydragnn.ops import radius_graph
    from hydragnn.utils import synthetic as syn

    spec = syn.WORKLOADS[name]
    batch = syn.collate(graphs)
    if spec["edges"] == "radius":
        g = radius_graph(batch["pos"].to(device), batch["ptr"], spec["radius"], spec["max_neighbours"],
                         loop=False, cell=batch["cell"] if spec["pbc"] else None,
                         pbc=batch["pbc"] if spec["pbc"] else None)
        batch["edge_index"] = g["edge_index"].cpu()
        if spec["pbc"]:
            batch["edge_shifts"] = g["edge_shifts"].cpu()
    return syn.finish_edges(batch)


def build_batch_cpu(name, graphs):
    """Same batch without touching the GPU (reference arm): scipy periodic KD-tree for the cubic
    crystals, the oracle's ASE / torch_cluster restatement otherwise."""
    from hydragnn.utils import synthetic as syn
    from oracle import kernel_spec as ks

    spec = syn.WORKLOADS[name]
    batch = syn.collate(graphs)
    if spec["edges"] == "radius":
        if name == "schnet_gps_fesi1024":
            from scipy.spatial import cKDTree
            srcs, dsts = [], []
            for g, (b, e) in enumerate(zip(batch["ptr"][:-1].tolist(), batch["ptr"][1:].tolist())):
                box = float(batch["cell"][g][0, 0])
                p = np.mod(batch["pos"][b:e].double().numpy(), box)
                p[p >= box] -= box
                pairs = cKDTree(p, boxsize=box).query_pairs(spec["radius"], output_type="ndarray")
                i = np.concatenate([pairs[:, 0], pairs[:, 1]]) + b
                j = np.concatenate([pairs[:, 1], pairs[:, 0]]) + b
                order = np.lexsort((i, j))
                srcs.append(i[order]); dsts.append(j[order])
            batch["edge_index"] = torch.from_numpy(np.stack([np.concatenate(srcs), np.concatenate(dsts)])).long()
            batch["edge_shifts"] = torch.zeros(batch["edge_index"].size(1), 3)   # dropped by SchNet (F7)
        else:
            src, dst, S, _, _ = ks.radius_graph(batch["pos"], batch["ptr"], spec["radius"], spec["max_neighbours"],
                                                False, batch.get("cell"), batch.get("pbc"), mode=0 if spec["pbc"] else 1)
            batch["edge_index"] = torch.stack([src, dst]).long()
            if spec["pbc"]:
                gid = torch.bucketize(dst.long(), batch["ptr"][1:], right=True)
                batch["edge_shifts"] = torch.einsum("ek,ekd->ed", S.float(), batch["cell"][gid])
    return syn.finish_edges(batch)


TENSOR_FIELDS = ("x", "pos", "pe", "rel_pe", "edge_index", "edge_attr", "edge_shifts", "batch", "ptr",
                 "y", "energy", "forces")


def batch_bytes(batch):
    return int(sum(batch[k].numel() * batch[k].element_size() for k in TENSOR_FIELDS if k in batch))


def to_data(batch, device, data_cls, non_blocking=False):
    fields = {k: batch[k].to(device, non_blocking=non_blocking) for k in TENSOR_FIELDS if k in batch}
    fields["num_graphs"] = batch["num_graphs"]
    return data_cls(**fields)
