"""Generate tests/golden/pbc_retry_fixtures.pt from the REAL reference RadiusGraphPBC.

Same mechanism as make_golden.py (the reference's own source file is imported unmodified, ASE is
replaced by oracle/pyg_lite.neighbor_list_ase).  The cases exercise the parts of
RadiusGraphPBC.__call__ the first fixture file does not reach:
  * the cutoff expansion x1.25 when a node receives no edge (up to 3 attempts,
    hydragnn/preprocess/graph_samples_checks_and_updates.py:163-206),
  * the artificial-edge fallback `_ensure_connected` (:284-307), which draws its source nodes from
    numpy's global RNG -- seeded here, and by the tests, with SEED immediately before each sample.

    python tests/golden/make_golden_pbc_retry.py
"""
import contextlib
import io
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402

SEED = 20240229


def cluster(gen, n, centre, spread):
    return torch.tensor(centre, dtype=torch.float32) + (torch.rand(n, 3, generator=gen) - 0.5) * spread


def cases():
    gen = torch.Generator().manual_seed(11)
    big = torch.eye(3) * 40.0
    out = {}
    # one expansion: a far atom 2.8 away from a tight cluster, r = 2.5 -> 3.125
    out["retry_once"] = dict(pos=torch.cat([cluster(gen, 5, [5.0, 5.0, 5.0], 1.6), torch.tensor([[8.6, 5.1, 4.9]])]),
                             cell=big, pbc=[False, False, False], r=2.5, cap=100000)
    # two expansions: r = 2.0 -> 2.5 -> 3.125
    out["retry_twice"] = dict(pos=torch.cat([cluster(gen, 6, [5.0, 5.0, 5.0], 1.6), torch.tensor([[8.55, 5.0, 5.0]])]),
                              cell=big, pbc=[False, False, False], r=2.0, cap=100000)
    # never connected: two atoms far away -> artificial edges after the third attempt
    out["artificial"] = dict(pos=torch.cat([cluster(gen, 5, [5.0, 5.0, 5.0], 1.6),
                                            torch.tensor([[20.0, 20.0, 20.0], [30.0, 8.0, 31.0]])]),
                             cell=big, pbc=[False, False, False], r=1.5, cap=100000)
    # periodic cell with a neighbour cap: the expansion changes which 4 neighbours are nearest for nobody,
    # but adds the edges of the far atom through the periodic image
    cell = torch.tensor([[9.0, 0.0, 0.0], [1.0, 8.0, 0.0], [0.0, 0.5, 30.0]])
    out["retry_pbc_cap"] = dict(pos=torch.cat([cluster(gen, 7, [2.0, 2.0, 2.0], 2.4), torch.tensor([[6.6, 5.9, 17.0]])]),
                                cell=cell, pbc=[True, True, False], r=4.0, cap=4)
    # a single atom in a non-periodic box: num_nodes == 1 branch of _ensure_connected (self edge 0 -> 0)
    out["single_atom"] = dict(pos=torch.tensor([[1.0, 2.0, 3.0]]), cell=big, pbc=[False, False, False],
                              r=1.0, cap=100000)
    # fully connected at the first attempt (control)
    out["control"] = dict(pos=cluster(gen, 9, [4.0, 4.0, 4.0], 3.0), cell=torch.eye(3) * 6.0,
                          pbc=[True, True, True], r=2.5, cap=100000)
    return out


def main():
    ref = mg.load_reference()
    fixtures = {}
    for name, c in cases().items():
        for loop in (False, True):
            d = mg.Data(pos=c["pos"].clone(), cell=c["cell"].clone(), pbc=list(c["pbc"]),
                        x=torch.zeros(c["pos"].size(0), 1))
            np.random.seed(SEED)
            log = io.StringIO()
            with contextlib.redirect_stdout(log):
                d = ref.graphs.RadiusGraphPBC(r=c["r"], loop=loop, max_num_neighbors=c["cap"])(d)
            fixtures[f"{name}_loop{int(loop)}"] = dict(
                pos=c["pos"], cell=c["cell"], pbc=torch.tensor(c["pbc"]), r=c["r"], cap=c["cap"], loop=loop,
                seed=SEED, expansions=log.getvalue().count("expanding radius"),
                artificial=log.getvalue().count("constructed artificially"),
                edge_index=d.edge_index, edge_shifts=d.edge_shifts)
            f = fixtures[f"{name}_loop{int(loop)}"]
            print(name, loop, "edges", d.edge_index.size(1), "expansions", f["expansions"],
                  "artificial", f["artificial"])
    path = os.path.join(HERE, "pbc_retry_fixtures.pt")
    torch.save(fixtures, path)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
