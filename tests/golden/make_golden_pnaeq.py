"""Generate tests/golden/pnaeq_fixtures.pt from the REAL reference PNAEqStack
(hydragnn/models/PNAEqStack.py, imported unmodified on top of oracle/pyg_lite.py).

    python tests/golden/make_golden_pnaeq.py
"""
import importlib
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402
import make_golden_pnaplus as mp  # noqa: E402
from oracle import pyg_lite as pyg  # noqa: E402

DEG = [0, 3, 6, 9, 8, 5, 3, 2, 1, 1, 1]
CASES = {
    "pnaeq_plain": dict(input_dim=3, hidden_dim=8, radius=3.0, max_neighbours=10, num_conv_layers=3, output_dim=[2],
                        output_type=["graph"], heads=mg.HEADS_GRAPH, task_weights=[1.0], num_graphs=4, n_lo=4,
                        n_hi=8, box=3.5, num_radial=5, pna_deg=DEG),
    "pnaeq_forces": dict(input_dim=3, hidden_dim=8, radius=3.0, max_neighbours=10, num_conv_layers=2, output_dim=[1],
                         output_type=["node"], heads=mg.HEADS_NODE, task_weights=[1.0], num_graphs=3, n_lo=5,
                         n_hi=8, box=3.5, num_nodes=8, num_radial=5, pna_deg=DEG, forces=True, pbc_shifts=True),
    "pnaeq_gps_edge": dict(input_dim=2, hidden_dim=8, radius=3.0, max_neighbours=10, num_conv_layers=2,
                           output_dim=[1], output_type=["graph"], heads=mg.HEADS_GRAPH, task_weights=[1.0],
                           num_graphs=4, n_lo=4, n_hi=8, box=3.5, edge_dim=1, num_radial=4, pna_deg=DEG, pe_dim=4,
                           engine="GPS", nheads=2, eval=True),
}


def install_extra_shims():
    mp.install_extra_shims()
    mg._module("torch_geometric.nn.aggr.scaler", DegreeScalerAggregation=pyg.DegreeScalerAggregation)
    nn_mod = sys.modules["torch_geometric.nn"]
    nn_mod.Linear = torch.nn.Linear
    sys.modules["torch_geometric"].nn = nn_mod


def build(cls, cfg):
    """Positional signature of create.py:446-470."""
    torch.manual_seed(0)
    kw = dict(loss_weights=cfg["task_weights"], freeze_conv=False, num_conv_layers=cfg["num_conv_layers"],
              num_nodes=cfg.get("num_nodes"))
    common = mg.common_args(cfg["input_dim"], cfg["hidden_dim"], cfg["output_dim"], cfg["output_type"], cfg["heads"],
                            cfg.get("pe_dim", 0), cfg.get("engine"), cfg.get("nheads", 0))
    sig = "inv_node_feat, equiv_node_feat, edge_index, edge_rbf, edge_vec"
    return cls(sig, sig, cfg["pna_deg"], cfg.get("edge_dim"), cfg["num_radial"], cfg["radius"], *common, **kw)


def main():
    mg.load_reference()
    install_extra_shims()
    stack = importlib.import_module("hydragnn.models.PNAEqStack").PNAEqStack
    fixtures = {}
    for name, cfg in CASES.items():
        gen = torch.Generator().manual_seed(987 + len(name))
        batch = mg.make_batch(gen, cfg)
        model = build(stack, cfg)
        if cfg.get("eval"):
            model.eval()
        out = mg.run_model(model, batch, cfg)
        fixtures[name] = dict(kind="PNAEq", cfg=cfg, batch=batch, out=out,
                              state={k: v.detach().clone() for k, v in model.state_dict().items()})
        print(f"{name}: loss={float(out['loss']):.6f} params={sum(p.numel() for p in model.parameters())}")
    path = os.path.join(HERE, "pnaeq_fixtures.pt")
    torch.save(fixtures, path)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
