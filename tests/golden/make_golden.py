"""Generate tests/golden/*.pt from the REAL reference modules.

Runs only in the development container (it reads /root/reference, which does
not exist on the GPU box).  The reference cannot be imported as is -- PyG,
torch_scatter, torch_cluster, ase and mpi4py are not installed -- so this
script registers oracle/pyg_lite.py (a restatement of those libraries'
semantics) under the missing module names, stubs the heavy package
__init__ files, and then imports the reference's own source files unmodified:

    hydragnn/models/{Base,SCFStack,EGCLStack,PAINNStack,PNAStack}.py
    hydragnn/globalAtt/gps.py
    hydragnn/utils/model/{operations,model}.py
    hydragnn/preprocess/graph_samples_checks_and_updates.py

Each fixture stores seeded inputs, the module's state_dict and the outputs the
reference code produced, so that tests/test_oracle_golden.py can check
oracle/hydragnn_ref.py (and, on the GPU, the product modules) against them.

    python tests/golden/make_golden.py
"""
import importlib
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, ROOT)

from oracle import pyg_lite as pyg  # noqa: E402


class Data:
    """Minimal torch_geometric.data.Data stand-in (attribute bag)."""

    def __init__(self, **kw):
        for k, v in kw.items():
            setattr(self, k, v)

    def __contains__(self, key):
        return hasattr(self, key) and getattr(self, key) is not None

    @property
    def num_nodes(self):
        return self.pos.size(0) if hasattr(self, "pos") else self.x.size(0)


def _module(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def install_shims():
    class RadiusGraph:
        def __init__(self, r, loop=False, max_num_neighbors=32, flow="source_to_target", num_workers=1):
            self.r, self.loop, self.max_num_neighbors = r, loop, max_num_neighbors

        def __call__(self, data):
            data.edge_index = pyg.radius_graph_cluster(
                data.pos, self.r, getattr(data, "batch", None), self.loop, self.max_num_neighbors)
            return data

    class Distance:
        def __init__(self, norm=True, max_value=None, cat=True):
            self.norm, self.max, self.cat = norm, max_value, cat

        def __call__(self, data):
            row, col = data.edge_index
            dist = torch.norm(data.pos[col] - data.pos[row], p=2, dim=-1).view(-1, 1)
            if self.norm and dist.numel() > 0:
                dist = dist / (dist.max() if self.max is None else self.max)
            prev = getattr(data, "edge_attr", None)
            data.edge_attr = dist if prev is None or not self.cat else torch.cat(
                [prev.view(-1, 1) if prev.dim() == 1 else prev, dist.type_as(prev)], dim=-1)
            return data

    class LocalCartesian(Distance):
        pass

    def functional_transform(name):
        return lambda cls: cls

    def remove_self_loops(edge_index, edge_attr=None):
        mask = edge_index[0] != edge_index[1]
        return edge_index[:, mask], (None if edge_attr is None else edge_attr[mask])

    class PerformerAttention(torch.nn.Module):
        def __init__(self, *a, **k):
            super().__init__()
            raise NotImplementedError

    schnet = _module("torch_geometric.nn.models.schnet", GaussianSmearing=pyg.GaussianSmearing,
                     RadiusInteractionGraph=pyg.RadiusInteractionGraph, ShiftedSoftplus=pyg.ShiftedSoftplus)
    models = _module("torch_geometric.nn.models", schnet=schnet)
    conv = _module("torch_geometric.nn.conv", MessagePassing=pyg.MessagePassing, PNAConv=pyg.PNAConv)
    attention = _module("torch_geometric.nn.attention", PerformerAttention=PerformerAttention)
    inits = _module("torch_geometric.nn.inits", reset=pyg.reset)
    resolver = _module("torch_geometric.nn.resolver", activation_resolver=pyg.activation_resolver,
                       normalization_resolver=pyg.normalization_resolver)
    nn_mod = _module("torch_geometric.nn", Sequential=pyg.Sequential, MessagePassing=pyg.MessagePassing,
                     PNAConv=pyg.PNAConv, BatchNorm=pyg.BatchNorm, global_mean_pool=pyg.global_mean_pool,
                     models=models, conv=conv, attention=attention, inits=inits, resolver=resolver)
    typing_mod = _module("torch_geometric.typing", OptTensor=torch.Tensor, Adj=torch.Tensor)
    utils_mod = _module("torch_geometric.utils", to_dense_batch=pyg.to_dense_batch, degree=pyg.degree,
                        remove_self_loops=remove_self_loops, scatter=pyg.scatter)
    transforms = _module("torch_geometric.transforms", RadiusGraph=RadiusGraph, Distance=Distance,
                         LocalCartesian=LocalCartesian)
    datapipes = _module("torch_geometric.data.datapipes", functional_transform=functional_transform)
    data_mod = _module("torch_geometric.data", Data=Data, datapipes=datapipes)
    _module("torch_geometric", nn=nn_mod, typing=typing_mod, utils=utils_mod, transforms=transforms,
            data=data_mod)
    _module("torch_scatter", scatter=pyg.scatter, scatter_add=pyg.scatter_add)

    class _Comm:
        def Barrier(self):
            pass

        def Get_rank(self):
            return 0

        def Get_size(self):
            return 1

    mpi = _module("mpi4py.MPI", COMM_WORLD=_Comm(), SUM=None, MAX=None)
    _module("mpi4py", MPI=mpi)

    class Atoms:
        def __init__(self, positions, cell, pbc):
            self.positions = np.asarray(positions, dtype=np.float64)
            self.cell = np.asarray(cell, dtype=np.float64)
            self.pbc = np.asarray(pbc, dtype=bool)

    def neighbor_list(quantities, a, cutoff, self_interaction=False):
        assert quantities == "ijdS"
        return pyg.neighbor_list_ase(a.positions, a.cell, a.pbc, cutoff, self_interaction)

    nl = _module("ase.neighborlist", neighbor_list=neighbor_list)
    _module("ase", Atoms=Atoms, neighborlist=nl)


def stub_packages():
    """Give the heavy reference packages a __path__ without running their __init__."""
    for name, sub in [("hydragnn", ""), ("hydragnn.models", "models"), ("hydragnn.utils", "utils"),
                      ("hydragnn.preprocess", "preprocess"), ("hydragnn.globalAtt", "globalAtt")]:
        m = types.ModuleType(name)
        m.__path__ = [os.path.join(REF, "hydragnn", sub)]
        sys.modules[name] = m
    # dataset_descriptors pulls in mendeleev-free code only; import it for real
    sys.modules["hydragnn.preprocess.dataset_descriptors"] = types.ModuleType(
        "hydragnn.preprocess.dataset_descriptors")
    sys.modules["hydragnn.preprocess.dataset_descriptors"].AtomFeatures = object


def load_reference():
    install_shims()
    stub_packages()
    ref = types.SimpleNamespace()
    ref.operations = importlib.import_module("hydragnn.utils.model.operations")
    ref.model_utils = importlib.import_module("hydragnn.utils.model.model")
    ref.gps = importlib.import_module("hydragnn.globalAtt.gps")
    ref.Base = importlib.import_module("hydragnn.models.Base")
    ref.SCF = importlib.import_module("hydragnn.models.SCFStack")
    ref.EGCL = importlib.import_module("hydragnn.models.EGCLStack")
    ref.PAINN = importlib.import_module("hydragnn.models.PAINNStack")
    ref.PNA = importlib.import_module("hydragnn.models.PNAStack")
    ref.graphs = importlib.import_module("hydragnn.preprocess.graph_samples_checks_and_updates")
    return ref


# ------------------------------------------------------------------ inputs ---
def random_molecules(gen, num_graphs, n_lo, n_hi, box, min_sep=0.8):
    pos, batch = [], []
    for g in range(num_graphs):
        n = int(torch.randint(n_lo, n_hi + 1, (1,), generator=gen))
        pts = []
        while len(pts) < n:
            p = torch.rand(3, generator=gen) * box
            if all(torch.norm(p - q) > min_sep for q in pts):
                pts.append(p)
        pos.append(torch.stack(pts))
        batch.append(torch.full((n,), g, dtype=torch.long))
    return torch.cat(pos), torch.cat(batch)


HEADS_GRAPH = {"graph": [{"type": "branch-0", "architecture": {
    "num_sharedlayers": 2, "dim_sharedlayers": 10, "num_headlayers": 2, "dim_headlayers": [12, 6]}}]}
HEADS_NODE = {"node": [{"type": "branch-0", "architecture": {
    "num_headlayers": 2, "dim_headlayers": [10, 10], "type": "mlp"}}]}
HEADS_MIXED = {"graph": HEADS_GRAPH["graph"], "node": HEADS_NODE["node"]}


def common_args(input_dim, hidden, output_dim, output_type, heads, pe_dim=0, engine=None, nheads=0,
                equivariance=False, loss="mse", act="relu"):
    return (input_dim, hidden, output_dim, pe_dim, engine, "multihead" if engine else None, nheads,
            output_type, heads, act, loss, equivariance)


def build(ref_or_oracle, kind, cfg):
    """Instantiate a stack through the same positional signature create.py uses (:135-425)."""
    M = ref_or_oracle
    torch.manual_seed(0)
    kw = dict(loss_weights=cfg["task_weights"], freeze_conv=False, initial_bias=None,
              num_conv_layers=cfg["num_conv_layers"], num_nodes=cfg.get("num_nodes"))
    if "dropout" in cfg:        # train-mode fixtures (make_golden_train.py): Base.__init__(dropout=...)
        kw["dropout"] = cfg["dropout"]
    common = common_args(cfg["input_dim"], cfg["hidden_dim"], cfg["output_dim"], cfg["output_type"],
                         cfg["heads"], cfg.get("pe_dim", 0), cfg.get("engine"), cfg.get("nheads", 0),
                         cfg.get("equivariance", False))
    if kind == "SchNet":
        return M["SCFStack"]("", "inv_node_feat, equiv_node_feat, edge_index, edge_weight, edge_rbf",
                             cfg["num_filters"], cfg.get("edge_dim"), cfg["num_gaussians"], cfg["radius"],
                             *common, max_neighbours=cfg["max_neighbours"], **kw)
    if kind == "EGNN":
        return M["EGCLStack"]("inv_node_feat, equiv_node_feat, edge_index, edge_attr", "",
                              cfg.get("edge_dim"), *common, max_neighbours=cfg["max_neighbours"], **kw)
    if kind == "PAINN":
        sig = "inv_node_feat, equiv_node_feat, edge_index, diff, dist"
        return M["PAINNStack"](sig, sig, cfg.get("edge_dim"), cfg["num_radial"], cfg["radius"], *common, **kw)
    if kind == "PNA":
        return M["PNAStack"]("inv_node_feat, equiv_node_feat, edge_index", "inv_node_feat, edge_index",
                             cfg["pna_deg"], cfg.get("edge_dim"), *common, **kw)
    raise ValueError(kind)


def ref_classes(ref):
    return {"SCFStack": ref.SCF.SCFStack, "EGCLStack": ref.EGCL.EGCLStack,
            "PAINNStack": ref.PAINN.PAINNStack, "PNAStack": ref.PNA.PNAStack}


def make_batch(gen, cfg):
    """Seeded synthetic mini-batch with every field the four stacks may read."""
    pos, batch = random_molecules(gen, cfg["num_graphs"], cfg["n_lo"], cfg["n_hi"], cfg["box"])
    N = pos.size(0)
    edge_index = pyg.radius_graph_cluster(pos, cfg["radius"], batch, False, cfg["max_neighbours"])
    E = edge_index.size(1)
    d = dict(pos=pos, batch=batch, edge_index=edge_index,
             x=torch.randn(N, cfg["input_dim"], generator=gen))
    if cfg.get("edge_dim"):
        d["edge_attr"] = torch.rand(E, cfg["edge_dim"], generator=gen)
    if cfg.get("pe_dim"):
        pe = torch.randn(N, cfg["pe_dim"], generator=gen) * 0.3
        d["pe"] = pe
        d["rel_pe"] = (pe[edge_index[0]] - pe[edge_index[1]]).abs()
    if cfg.get("pbc_shifts"):
        d["edge_shifts"] = torch.randint(-1, 2, (E, 3), generator=gen).float() * 0.05
    d["energy"] = torch.randn(cfg["num_graphs"], generator=gen)
    d["forces"] = torch.randn(N, 3, generator=gen) * 0.1
    return d


def run_model(model, d, cfg):
    """Forward + the loss the config asks for + parameter gradients."""
    data = Data(**{k: (v.clone() if torch.is_tensor(v) else v) for k, v in d.items()})
    out = {}
    if cfg.get("forces"):
        data.pos.requires_grad_(True)
        pred = model(data)
        loss, tasks = model.energy_force_loss(pred, data)[:2]
        out["pred0"] = pred[0].detach()
    else:
        pred = model(data)
        gen = torch.Generator().manual_seed(99)
        ys, head_index, off = [], [], 0
        for p in pred:
            ys.append(torch.randn(p.numel(), 1, generator=gen))
            head_index.append(torch.arange(off, off + p.numel()))
            off += p.numel()
        y = torch.cat(ys)
        out["y"], out["head_index"] = y, head_index
        loss, tasks = model.loss(pred, y, head_index)
        for k, p in enumerate(pred):
            out[f"pred{k}"] = p.detach()
    loss.backward()
    out["loss"] = loss.detach()
    out["grads"] = {n: p.grad.detach().clone() for n, p in model.named_parameters() if p.grad is not None}
    if cfg.get("forces"):
        out["pos_grad"] = data.pos.grad.detach().clone()
    return out


MODEL_CASES = {
    "schnet_plain": ("SchNet", dict(input_dim=3, hidden_dim=16, num_filters=16, num_gaussians=12, radius=3.0,
                                    max_neighbours=8, num_conv_layers=3, output_dim=[2], output_type=["graph"],
                                    heads=HEADS_GRAPH, task_weights=[1.0], num_graphs=5, n_lo=4, n_hi=9, box=4.0)),
    "schnet_equiv": ("SchNet", dict(input_dim=3, hidden_dim=16, num_filters=16, num_gaussians=12, radius=3.0,
                                    max_neighbours=8, num_conv_layers=3, output_dim=[2], output_type=["graph"],
                                    heads=HEADS_GRAPH, task_weights=[1.0], num_graphs=4, n_lo=4, n_hi=8, box=4.0,
                                    equivariance=True)),
    "schnet_gps": ("SchNet", dict(input_dim=3, hidden_dim=16, num_filters=16, num_gaussians=12, radius=3.0,
                                  max_neighbours=8, num_conv_layers=2, output_dim=[1], output_type=["graph"],
                                  heads=HEADS_GRAPH, task_weights=[1.0], num_graphs=4, n_lo=4, n_hi=8, box=4.0,
                                  pe_dim=4, engine="GPS", nheads=4, eval=True)),
    "egnn_gps_edge": ("EGNN", dict(input_dim=2, hidden_dim=16, radius=2.5, max_neighbours=6, num_conv_layers=3,
                                   output_dim=[1, 2], output_type=["graph", "node"], heads=HEADS_MIXED,
                                   task_weights=[1.0, 2.0], num_graphs=4, n_lo=5, n_hi=9, box=4.0, edge_dim=1,
                                   pe_dim=4, engine="GPS", nheads=2, num_nodes=9, eval=True)),
    "egnn_equiv": ("EGNN", dict(input_dim=2, hidden_dim=12, radius=2.5, max_neighbours=6, num_conv_layers=3,
                                output_dim=[1], output_type=["graph"], heads=HEADS_GRAPH, task_weights=[1.0],
                                num_graphs=4, n_lo=5, n_hi=9, box=4.0, equivariance=True)),
    "painn_forces": ("PAINN", dict(input_dim=4, hidden_dim=12, radius=3.0, num_radial=5, max_neighbours=10,
                                   num_conv_layers=3, output_dim=[1], output_type=["node"], heads=HEADS_NODE,
                                   task_weights=[1.0], num_graphs=3, n_lo=5, n_hi=8, box=3.5, num_nodes=8,
                                   forces=True, pbc_shifts=True)),
    "schnet_forces": ("SchNet", dict(input_dim=3, hidden_dim=12, num_filters=12, num_gaussians=10, radius=3.0,
                                     max_neighbours=10, num_conv_layers=2, output_dim=[1], output_type=["node"],
                                     heads=HEADS_NODE, task_weights=[1.0], num_graphs=3, n_lo=5, n_hi=8, box=3.5,
                                     num_nodes=8, forces=True, edge_dim=1)),
    "egnn_forces": ("EGNN", dict(input_dim=3, hidden_dim=12, radius=3.0, max_neighbours=10, num_conv_layers=2,
                                 output_dim=[1], output_type=["node"], heads=HEADS_NODE, task_weights=[1.0],
                                 num_graphs=3, n_lo=5, n_hi=8, box=3.5, num_nodes=8, forces=True)),
    "pna_multihead": ("PNA", dict(input_dim=1, hidden_dim=8, radius=2.5, max_neighbours=8, num_conv_layers=2,
                                  output_dim=[1, 1, 1], output_type=["graph", "node", "node"], heads=HEADS_MIXED,
                                  task_weights=[1.0, 1.0, 1.0], num_graphs=6, n_lo=2, n_hi=8, box=3.0,
                                  num_nodes=8, pna_deg=[0, 3, 6, 9, 8, 5, 3, 2, 1])),
    "pna_edge": ("PNA", dict(input_dim=2, hidden_dim=8, radius=2.5, max_neighbours=8, num_conv_layers=2,
                             output_dim=[2], output_type=["graph"], heads=HEADS_GRAPH, task_weights=[1.0],
                             num_graphs=5, n_lo=3, n_hi=8, box=3.0, edge_dim=1,
                             pna_deg=[0, 3, 6, 9, 8, 5, 3, 2, 1])),
}


def main():
    ref = load_reference()
    classes = ref_classes(ref)
    fixtures = {}

    # (a) small functions -------------------------------------------------------
    gen = torch.Generator().manual_seed(1)
    pos = torch.randn(7, 3, generator=gen)
    ei = torch.randint(0, 7, (2, 15), generator=gen)
    sh = torch.randn(15, 3, generator=gen) * 0.1
    v1, l1 = ref.operations.get_edge_vectors_and_lengths(pos, ei, sh)
    v2, l2 = ref.operations.get_edge_vectors_and_lengths(pos, ei, sh, normalize=True, eps=1.0)
    data = torch.randn(15, 3, generator=gen)
    dist = torch.rand(15, 1, generator=gen) * 4 + 0.2
    fixtures["functions"] = dict(
        pos=pos, edge_index=ei, shifts=sh, vec=v1, len=l1, vec_norm=v2, len_norm=l2, data=data,
        seg_mean=ref.model_utils.unsorted_segment_mean(data, ei[0], 7),
        seg_sum=ref.EGCL.unsorted_segment_sum(data, ei[0], 7), dist=dist,
        sinc=ref.PAINN.sinc_expansion(dist, 5, 3.0), cutoff=ref.PAINN.cosine_cutoff(dist, 3.0))

    # (b) full models -----------------------------------------------------------
    for name, (kind, cfg) in MODEL_CASES.items():
        gen = torch.Generator().manual_seed(1234 + len(name))
        batch = make_batch(gen, cfg)
        model = build(classes, kind, cfg)
        if cfg.get("eval"):
            model.eval()       # GPS dropout (p=0.25) is not reproducible across devices
        out = run_model(model, batch, cfg)
        fixtures[name] = dict(kind=kind, cfg=cfg, batch=batch, out=out,
                              state={k: v.detach().clone() for k, v in model.state_dict().items()})
        print(f"{name}: loss={float(out['loss']):.6f} params={sum(p.numel() for p in model.parameters())}")

    # (c) periodic radius graphs (tests/test_periodic_boundary_conditions.py:24-127) ----
    pbc_cases = {}
    h2 = Data(pos=torch.tensor([[1.43, 0.0, 0.0], [2.17, 0.0, 0.0]]) * 1.0,
              cell=torch.eye(3) * 3.0, pbc=[True, True, True], x=torch.zeros(2, 1))
    bcc = []
    for i in range(5):
        for j in range(5):
            for k in range(5):
                bcc.append([i, j, k])
                bcc.append([i + 0.5, j + 0.5, k + 0.5])
    cr = Data(pos=torch.tensor(bcc, dtype=torch.float32) * 3.6, cell=torch.eye(3) * 18.0,
              pbc=[True, True, True], x=torch.zeros(250, 1))
    gen = torch.Generator().manual_seed(5)
    tri_cell = torch.tensor([[6.0, 0.0, 0.0], [1.5, 5.5, 0.0], [0.8, 1.1, 7.0]])
    tri = Data(pos=(torch.rand(40, 3, generator=gen) * 2 - 0.5) @ tri_cell, cell=tri_cell,
               pbc=[True, True, False], x=torch.zeros(40, 1))
    for cname, d, r, cap in [("h2", h2, 0.9, 100000), ("cr", cr, 5.0, 100000), ("tri", tri, 3.2, 100000),
                             ("tri_cap", tri, 3.2, 6), ("h2_big", h2, 3.5, 100000)]:
        for loop in (False, True):
            dd = Data(pos=d.pos.clone(), cell=d.cell.clone(), pbc=list(d.pbc), x=d.x)
            dd = ref.graphs.RadiusGraphPBC(r=r, loop=loop, max_num_neighbors=cap)(dd)
            pbc_cases[f"{cname}_loop{int(loop)}"] = dict(
                pos=d.pos, cell=d.cell, pbc=torch.tensor(d.pbc), r=r, cap=cap, loop=loop,
                edge_index=dd.edge_index, edge_shifts=dd.edge_shifts)
            print(cname, loop, "edges/atom", dd.edge_index.size(1) / d.pos.size(0))
    fixtures["pbc"] = pbc_cases

    torch.save(fixtures, os.path.join(HERE, "reference_fixtures.pt"))
    print("wrote", os.path.join(HERE, "reference_fixtures.pt"),
          os.path.getsize(os.path.join(HERE, "reference_fixtures.pt")) // 1024, "KiB")


if __name__ == "__main__":
    main()
