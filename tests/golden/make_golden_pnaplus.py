"""Generate tests/golden/pnaplus_fixtures.pt from the REAL reference PNAPlusStack
(hydragnn/models/PNAPlusStack.py, imported unmodified; PyG pieces it needs -- MessagePassing with a
DegreeScalerAggregation module, dense Linear, BesselBasisLayer -- come from oracle/pyg_lite.py).

    python tests/golden/make_golden_pnaplus.py
"""
import importlib
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402
from oracle import pyg_lite as pyg  # noqa: E402

CASES = {
    "pnaplus_plain": dict(input_dim=2, hidden_dim=8, radius=2.5, max_neighbours=8, num_conv_layers=2, output_dim=[2],
                          output_type=["graph"], heads=mg.HEADS_GRAPH, task_weights=[1.0], num_graphs=5, n_lo=3,
                          n_hi=8, box=3.0, num_radial=5, envelope_exponent=5,
                          pna_deg=[0, 3, 6, 9, 8, 5, 3, 2, 1]),
    "pnaplus_edge": dict(input_dim=2, hidden_dim=8, radius=2.5, max_neighbours=8, num_conv_layers=2, output_dim=[1, 1],
                         output_type=["graph", "node"], heads=mg.HEADS_MIXED, task_weights=[1.0, 0.5],
                         num_graphs=5, n_lo=3, n_hi=8, box=3.0, num_nodes=8, edge_dim=1, num_radial=6,
                         envelope_exponent=5, pna_deg=[0, 3, 6, 9, 8, 5, 3, 2, 1]),
    "pnaplus_gps_edge": dict(input_dim=2, hidden_dim=8, radius=2.5, max_neighbours=8, num_conv_layers=2,
                             output_dim=[1], output_type=["graph"], heads=mg.HEADS_GRAPH, task_weights=[1.0],
                             num_graphs=4, n_lo=4, n_hi=8, box=3.0, edge_dim=1, num_radial=5, envelope_exponent=5,
                             pna_deg=[0, 3, 6, 9, 8, 5, 3, 2, 1], pe_dim=4, engine="GPS", nheads=2, eval=True),
}


def install_extra_shims():
    mg._module("torch_geometric.nn.aggr", DegreeScalerAggregation=pyg.DegreeScalerAggregation)
    mg._module("torch_geometric.nn.dense.linear", Linear=torch.nn.Linear)
    mg._module("torch_geometric.nn.dense", linear=sys.modules["torch_geometric.nn.dense.linear"])
    mg._module("torch_geometric.nn.models.dimenet", BesselBasisLayer=pyg.BesselBasisLayer)


def build(cls, cfg):
    """Positional signature of create.py:192-214."""
    torch.manual_seed(0)
    kw = dict(loss_weights=cfg["task_weights"], freeze_conv=False, initial_bias=None,
              num_conv_layers=cfg["num_conv_layers"], num_nodes=cfg.get("num_nodes"))
    common = mg.common_args(cfg["input_dim"], cfg["hidden_dim"], cfg["output_dim"], cfg["output_type"], cfg["heads"],
                            cfg.get("pe_dim", 0), cfg.get("engine"), cfg.get("nheads", 0))
    return cls("inv_node_feat, equiv_node_feat, edge_index, rbf", "inv_node_feat, edge_index, rbf",
               cfg["pna_deg"], cfg.get("edge_dim"), cfg["envelope_exponent"], cfg["num_radial"], cfg["radius"],
               *common, **kw)


def main():
    ref = mg.load_reference()
    install_extra_shims()
    stack = importlib.import_module("hydragnn.models.PNAPlusStack").PNAPlusStack
    fixtures = {}
    for name, cfg in CASES.items():
        gen = torch.Generator().manual_seed(4321 + len(name))
        batch = mg.make_batch(gen, cfg)
        model = build(stack, cfg)
        if cfg.get("eval"):
            model.eval()
        out = mg.run_model(model, batch, cfg)
        fixtures[name] = dict(kind="PNAPlus", cfg=cfg, batch=batch, out=out,
                              state={k: v.detach().clone() for k, v in model.state_dict().items()})
        print(f"{name}: loss={float(out['loss']):.6f} params={sum(p.numel() for p in model.parameters())}",
              sorted(model.state_dict())[:4])
    path = os.path.join(HERE, "pnaplus_fixtures.pt")
    torch.save(fixtures, path)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
