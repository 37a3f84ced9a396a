"""Seeded synthetic workloads of the shapes BASELINE.json names.

No dataset can be downloaded here, so each workload draws atoms / molecules with the
published size statistics of the dataset it stands for.  A workload is
  * `arch`:   keyword arguments of hydragnn.models.create_model (random-init weights,
              torch.manual_seed(0) inside create_model as in the reference), and
  * `sample`: a function(rng) -> one graph as a dict of CPU tensors WITHOUT edges.
Edges are added by the caller's graph builder (the product uses the cell-list kernel,
the CPU baseline its own neighbour search), so both arms see identical atoms.
"""
import math

import numpy as np
import torch

GRAPH_HEAD = {"graph": [{"type": "branch-0", "architecture": {
    "num_sharedlayers": 2, "dim_sharedlayers": 50, "num_headlayers": 2, "dim_headlayers": [50, 25]}}]}
NODE_HEAD = {"node": [{"type": "branch-0", "architecture": {
    "num_headlayers": 2, "dim_headlayers": [200, 200], "type": "mlp"}}]}


def _arch(mpnn_type, **kw):
    base = dict(mpnn_type=mpnn_type, input_dim=1, hidden_dim=128, output_dim=[1], pe_dim=0,
                global_attn_engine=None, global_attn_type=None, global_attn_heads=0,
                output_type=["graph"], output_heads=GRAPH_HEAD, activation_function="relu",
                loss_function_type="mse", task_weights=[1.0], num_conv_layers=4,
                freeze_conv=False, initial_bias=None, num_nodes=None, max_neighbours=None,
                edge_dim=None, pna_deg=None, num_radial=None, num_gaussians=None, num_filters=None,
                radius=None, equivariance=False)
    base.update(kw)
    return base


# ------------------------------------------------------------------ samplers ---
def _qm9_molecule(rng):
    n = int(np.clip(round(rng.normal(18, 3)), 3, 29))
    side = (n / 0.1) ** (1.0 / 3.0)                       # 0.1 atoms per cubic Angstrom
    pts = []
    while len(pts) < n:
        p = rng.uniform(0, side, 3)
        if all(np.linalg.norm(p - q) > 0.9 for q in pts):
            pts.append(p)
    z = rng.choice([1, 6, 7, 8, 9], size=n, p=[0.51, 0.35, 0.06, 0.077, 0.003])
    return dict(pos=torch.tensor(np.array(pts), dtype=torch.float32),
                x=torch.tensor(z, dtype=torch.float32).view(-1, 1),
                y=torch.tensor(rng.normal(size=(1, 1)), dtype=torch.float32))


def _zinc_molecule(rng):
    """Random tree with unit bonds + ring closures: about 2.15 directed edges per atom."""
    n = int(np.clip(round(rng.normal(23.2, 4.5)), 9, 38))
    pos = np.zeros((n, 3))
    bonds = []
    for a in range(1, n):
        parent = int(rng.integers(max(0, a - 4), a))
        step = rng.normal(size=3)
        pos[a] = pos[parent] + step / np.linalg.norm(step)
        bonds.append((parent, a))
    for _ in range(max(0, int(round(0.075 * n)))):         # ring closures
        a, b = rng.integers(0, n, 2)
        if a != b and (min(a, b), max(a, b)) not in bonds:
            bonds.append((int(min(a, b)), int(max(a, b))))
    src = np.array([b[0] for b in bonds] + [b[1] for b in bonds])
    dst = np.array([b[1] for b in bonds] + [b[0] for b in bonds])
    kind = rng.integers(1, 4, len(bonds))
    order = np.lexsort((src, dst))
    pe = rng.normal(scale=1.0 / math.sqrt(n), size=(n, 8))
    return dict(pos=torch.tensor(pos, dtype=torch.float32),
                x=torch.tensor(rng.integers(0, 28, n), dtype=torch.float32).view(-1, 1),
                pe=torch.tensor(pe, dtype=torch.float32),
                edge_index=torch.tensor(np.stack([src, dst])[:, order], dtype=torch.long),
                edge_attr=torch.tensor(np.concatenate([kind, kind])[order], dtype=torch.float32).view(-1, 1),
                y=torch.tensor(rng.normal(size=(1, 1)), dtype=torch.float32))


def _oc20_slab(rng):
    n = int(np.clip(round(rng.normal(80, 15)), 40, 120))
    a, b = rng.uniform(10, 12, 2)
    gamma = math.radians(rng.uniform(60, 90))
    c = rng.uniform(30, 35)
    cell = np.array([[a, 0, 0], [b * math.cos(gamma), b * math.sin(gamma), 0], [0, 0, c]])
    n_ads = int(rng.integers(3, 9))
    frac = rng.uniform(0, 1, (n, 3))
    frac[:, 2] = rng.uniform(0, 12.0 / c, n)
    frac[n - n_ads:, 2] = rng.uniform(12.0 / c, 15.0 / c, n_ads)
    return dict(pos=torch.tensor(frac @ cell, dtype=torch.float32),
                x=torch.tensor(rng.integers(1, 84, n), dtype=torch.float32).view(-1, 1) / 83.0,
                cell=torch.tensor(cell, dtype=torch.float32), pbc=torch.tensor([True, True, True]),
                energy=torch.tensor(rng.normal(size=1), dtype=torch.float32),
                forces=torch.tensor(rng.normal(scale=0.05, size=(n, 3)), dtype=torch.float32))


def _fesi_crystal(rng, cells=8, a=2.80):
    grid = np.stack(np.meshgrid(*[np.arange(cells)] * 3, indexing="ij"), -1).reshape(-1, 3).astype(np.float64)
    pos = np.concatenate([grid, grid + 0.5]) * a
    pos = pos + rng.normal(scale=0.05, size=pos.shape)     # jitter breaks the lattice's distance ties
    n = pos.shape[0]
    species = rng.integers(0, 2, n)                        # Fe / Si with p = 0.5
    pe = rng.normal(scale=1.0 / math.sqrt(n), size=(n, 8))
    return dict(pos=torch.tensor(pos, dtype=torch.float32),
                x=torch.tensor(species, dtype=torch.float32).view(-1, 1),
                pe=torch.tensor(pe, dtype=torch.float32),
                cell=torch.eye(3) * (a * cells), pbc=torch.tensor([True, True, True]),
                y=torch.tensor(rng.normal(size=(1, 1)), dtype=torch.float32))


WORKLOADS = {
    # BASELINE.json configs[1]
    "schnet_qm9": dict(
        arch=_arch("SchNet", num_gaussians=50, num_filters=128, radius=5.0, max_neighbours=32),
        sample=_qm9_molecule, graphs_per_gpu=64, radius=5.0, max_neighbours=32, pbc=False,
        edges="radius", forces=False, seed=1236),
    # configs[2]
    "egnn_gps_zinc": dict(
        arch=_arch("EGNN", pe_dim=8, global_attn_engine="GPS", global_attn_type="multihead",
                   global_attn_heads=4, edge_dim=1),
        sample=_zinc_molecule, graphs_per_gpu=64, radius=None, max_neighbours=None, pbc=False,
        edges="given", forces=False, seed=1237),
    # configs[3]
    "painn_oc20_forces": dict(
        arch=_arch("PAINN", num_conv_layers=3, num_radial=6, radius=6.0, output_type=["node"],
                   output_heads=NODE_HEAD, num_nodes=120),
        sample=_oc20_slab, graphs_per_gpu=32, radius=6.0, max_neighbours=50, pbc=True,
        edges="radius", forces=True, seed=1238),
    # the model of the reference's physics-informed multibranch configuration
    # (PNAEq + force loss) on the OC20-shaped batch of configs[3]; degree histogram of a 50-neighbour cap
    "pnaeq_oc20_forces": dict(
        arch=_arch("PNAEq", num_conv_layers=3, num_radial=6, radius=6.0, output_type=["node"],
                   output_heads=NODE_HEAD, num_nodes=120,
                   pna_deg=[0] * 20 + [1, 2, 4, 8, 16, 32, 64, 128, 256, 512] + [1024] * 21),
        sample=_oc20_slab, graphs_per_gpu=32, radius=6.0, max_neighbours=50, pbc=True,
        edges="radius", forces=True, seed=1238),
    # configs[4] -- the multi-GPU headline
    "schnet_gps_fesi1024": dict(
        arch=_arch("SchNet", num_gaussians=50, num_filters=128, radius=5.0, max_neighbours=64,
                   pe_dim=8, global_attn_engine="GPS", global_attn_type="multihead", global_attn_heads=4),
        sample=_fesi_crystal, graphs_per_gpu=8, radius=5.0, max_neighbours=64, pbc=True,
        edges="radius", forces=False, seed=1239),
}


def sample_graphs(name, num_graphs, rank=0):
    """`num_graphs` seeded graphs of workload `name` (rank r draws its own shard)."""
    spec = WORKLOADS[name]
    rng = np.random.default_rng(spec["seed"] + 1000 * rank)
    return [spec["sample"](rng) for _ in range(num_graphs)]


def collate(graphs):
    """Concatenate edge-less graphs into one dict of CPU tensors with `batch` / `ptr`."""
    sizes = [g["pos"].size(0) for g in graphs]
    ptr = torch.tensor([0] + list(np.cumsum(sizes)), dtype=torch.long)
    out = dict(ptr=ptr, batch=torch.repeat_interleave(torch.arange(len(graphs)), torch.tensor(sizes)),
               num_graphs=len(graphs))
    for key in ("pos", "x", "pe", "forces"):
        if key in graphs[0]:
            out[key] = torch.cat([g[key] for g in graphs])
    for key in ("y",):
        if key in graphs[0]:
            out[key] = torch.cat([g[key] for g in graphs])
    if "energy" in graphs[0]:
        out["energy"] = torch.cat([g["energy"] for g in graphs])
    if "cell" in graphs[0]:
        out["cell"] = torch.stack([g["cell"] for g in graphs])
        out["pbc"] = torch.stack([g["pbc"] for g in graphs])
    if "edge_index" in graphs[0]:
        out["edge_index"] = torch.cat([g["edge_index"] + int(o) for g, o in zip(graphs, ptr[:-1])], dim=1)
        out["edge_attr"] = torch.cat([g["edge_attr"] for g in graphs])
    return out


def finish_edges(batch):
    """rel_pe = |pe_src - pe_dst| once the edge list exists (serialized_dataset_loader.py:183-189)."""
    if "pe" in batch and "edge_index" in batch:
        ei = batch["edge_index"]
        batch["rel_pe"] = (batch["pe"][ei[0]] - batch["pe"][ei[1]]).abs()
    return batch
