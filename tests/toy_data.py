"""Deterministic toy dataset in the reference's "unit_test" text format (what its CI trains on,
tests/deterministic_graph_data.py): BCC cells of 1-2 x 1-2 x 1 unit cells (2, 4 or 8 atoms), an integer
node type, nodal outputs x (kNN-smoothed type), x^2 + type, x^3 and their sum as the graph output.
Own implementation; only the file layout is shared with the reference."""
import os

import numpy as np
from sklearn.neighbors import KNeighborsRegressor


def write_configuration(path, index, cells, rng, num_types=3, num_neighbors=2):
    ux, uy, uz = cells
    base = np.array([[x, y, z] for x in range(ux) for y in range(uy) for z in range(uz)], dtype=np.float64)
    pos = np.empty((2 * len(base), 3))
    pos[0::2], pos[1::2] = base, base + 0.5
    n = pos.shape[0]
    kind = rng.integers(0, num_types, size=(n, 1)).astype(np.float64)
    smooth = KNeighborsRegressor(num_neighbors).fit(pos, kind).predict(pos)
    out1, out2, out3 = smooth, smooth ** 2 + kind, smooth ** 3
    total, total_linear = float(out1.sum() + out2.sum() + out3.sum()), float(out1.sum())
    rows = [f"{total:.6g}\t{total_linear:.6g}"]
    for i in range(n):
        vals = [kind[i, 0], float(i), *pos[i], out1[i, 0], out2[i, 0], out3[i, 0]]
        rows.append("\t".join(f"{v:.2f}" for v in vals))
    with open(os.path.join(path, f"output{index}.txt"), "w") as fh:
        fh.write("\n".join(rows))


def generate(path, count, seed, start=0):
    os.makedirs(path, exist_ok=True)
    rng = np.random.default_rng(seed)
    for i in range(count):
        cells = (int(rng.integers(1, 3)), int(rng.integers(1, 3)), 1)
        write_configuration(path, start + i, cells, rng)
