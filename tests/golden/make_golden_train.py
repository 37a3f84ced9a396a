"""Generate tests/golden/gps_train_fixtures.pt: GPS models of the REAL reference in TRAIN mode.

The fixtures of make_golden.py run the GPS cases in eval() because the reference's dropout
(p = 0.25, hydragnn/models/Base.py:53) is not reproducible across devices.  In eval the three
BatchNorms of every GPSConv (hydragnn/globalAtt/gps.py:80-83,116-150) normalise with their
initial running statistics, so the batch-statistics forward / backward -- the mode training and
bench.py run in -- was never compared with the reference.  Here the reference stacks are built
with `dropout=0.0` (Base.__init__ takes it, Base.py:53,67-68) and left in train(): BatchNorm uses
batch statistics and updates its running buffers, nothing is random.

Each case stores the seeded batch, the state dict BEFORE the step, loss / predictions /
parameter gradients, and the BatchNorm buffers AFTER the step.  One case has the BASELINE
widths (hidden 128, 50 gaussians, 4 heads) so that the tensor-core kernels are the ones checked.

    python tests/golden/make_golden_train.py        (development container only: reads /root/reference)
"""
import os

import torch

import make_golden as mg

HERE = os.path.dirname(os.path.abspath(__file__))

TRAIN_CASES = {
    "schnet_gps_train": ("SchNet", dict(
        input_dim=3, hidden_dim=16, num_filters=16, num_gaussians=12, radius=3.0, max_neighbours=8,
        num_conv_layers=2, output_dim=[1], output_type=["graph"], heads=mg.HEADS_GRAPH, task_weights=[1.0],
        num_graphs=4, n_lo=4, n_hi=8, box=4.0, pe_dim=4, engine="GPS", nheads=4, dropout=0.0)),
    "egnn_gps_edge_train": ("EGNN", dict(
        input_dim=2, hidden_dim=16, radius=2.5, max_neighbours=6, num_conv_layers=3, output_dim=[1, 2],
        output_type=["graph", "node"], heads=mg.HEADS_MIXED, task_weights=[1.0, 2.0], num_graphs=4, n_lo=5,
        n_hi=9, box=4.0, edge_dim=1, pe_dim=4, engine="GPS", nheads=2, num_nodes=9, dropout=0.0)),
    "painn_gps_train": ("PAINN", dict(
        input_dim=4, hidden_dim=12, radius=3.0, num_radial=5, max_neighbours=10, num_conv_layers=2,
        output_dim=[1], output_type=["graph"], heads=mg.HEADS_GRAPH, task_weights=[1.0], num_graphs=3,
        n_lo=5, n_hi=8, box=3.5, pe_dim=4, engine="GPS", nheads=2, dropout=0.0)),
    # BASELINE widths (config 5 / config 3 shapes): fused filter network, tcgen05 attention, native tail
    "schnet_gps_train_wide": ("SchNet", dict(
        input_dim=1, hidden_dim=128, num_filters=128, num_gaussians=50, radius=3.0, max_neighbours=16,
        num_conv_layers=2, output_dim=[1], output_type=["graph"], heads=mg.HEADS_GRAPH, task_weights=[1.0],
        num_graphs=3, n_lo=40, n_hi=60, box=6.5, pe_dim=8, engine="GPS", nheads=4, dropout=0.0)),
    "egnn_gps_train_wide": ("EGNN", dict(
        input_dim=1, hidden_dim=128, radius=2.5, max_neighbours=8, num_conv_layers=2, output_dim=[1],
        output_type=["graph"], heads=mg.HEADS_GRAPH, task_weights=[1.0], num_graphs=3, n_lo=30, n_hi=50,
        box=6.0, edge_dim=1, pe_dim=8, engine="GPS", nheads=4, dropout=0.0)),
}


def main():
    ref = mg.load_reference()
    classes = mg.ref_classes(ref)
    fixtures = {}
    for name, (kind, cfg) in TRAIN_CASES.items():
        gen = torch.Generator().manual_seed(4321 + len(name))
        batch = mg.make_batch(gen, cfg)
        model = mg.build(classes, kind, cfg)
        model.train()
        state = {k: v.detach().clone() for k, v in model.state_dict().items()}
        out = mg.run_model(model, batch, cfg)
        out["buffers"] = {k: v.detach().clone() for k, v in model.named_buffers()
                          if k.endswith(("running_mean", "running_var"))}
        fixtures[name] = dict(kind=kind, cfg=cfg, batch=batch, out=out, state=state, train=True)
        print(f"{name}: loss={float(out['loss']):.6f} N={batch['pos'].size(0)} E={batch['edge_index'].size(1)} "
              f"params={sum(p.numel() for p in model.parameters())}")
    path = os.path.join(HERE, "gps_train_fixtures.pt")
    torch.save(fixtures, path)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
