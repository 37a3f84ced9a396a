"""Dataset loading for the formats the CI configs use ("unit_test" / "LSMS" text files, or the
pickles this module writes), splitting, and DataLoaders.

Same entry point and on-disk layout as the reference (`dataset_loading_and_splitting(config)`,
preprocess/load_data.py:206-222; `$SERIALIZED_DATA_PATH/serialized_dataset/<name>[_split].pkl`
holding node min/max, graph min/max and the list of samples, raw_dataset_loader.py:89-163).
What differs is where the graph work happens: the reference builds one radius graph per sample on
the CPU in a Python loop (serialized_dataset_loader.py:134-150); here ALL samples of a split are
concatenated and handed to the cell-list kernel in one call (`ops.radius_graph` over the sample
`ptr`), and edge lengths / relative encodings are computed on the device for the whole split.
"""
import os
import pickle
import random

import numpy as np
import torch

from ..graph import Batch, Data
from ..utils.distributed import get_comm_size_and_rank, get_device


# ------------------------------------------------------------------ raw text files ---
def _parse_lsms_file(path, node_dims, node_cols, graph_dims, graph_cols):
    """First line: graph-level values; then one line per atom: feature, id, x, y, z, outputs...
    (reference: preprocess/lsms_raw_dataset_loader.py:50-93)."""
    with open(path, "r", encoding="utf-8") as fh:
        lines = [ln for ln in fh.read().splitlines() if ln.strip()]
    head = lines[0].split()
    y = [float(head[c + k]) for d, c in zip(graph_dims, graph_cols) for k in range(d)]
    pos, feats = [], []
    for line in lines[1:]:
        cols = line.split()
        pos.append([float(cols[2]), float(cols[3]), float(cols[4])])
        feats.append([float(cols[c + k]) for d, c in zip(node_dims, node_cols) for k in range(d)])
    x = torch.tensor(feats, dtype=torch.float32)
    if x.size(1) > 1:
        x[:, 1] = x[:, 1] - x[:, 0]          # LSMS charge density = column 1 - number of protons
    return Data(x=x, pos=torch.tensor(pos, dtype=torch.float32), y=torch.tensor(y, dtype=torch.float32))


def _minmax_normalize(splits, node_dims, graph_dims):
    """Min-max scale every node / graph feature over all splits (raw_dataset_loader.py:193-260)."""
    def ranges(dims):
        out, start = [], 0
        for d in dims:
            out.append((start, start + d))
            start += d
        return out

    node_rng, graph_rng = ranges(node_dims), ranges(graph_dims)
    node_mm = np.stack([np.full(len(node_dims), np.inf), np.full(len(node_dims), -np.inf)])
    graph_mm = np.stack([np.full(len(graph_dims), np.inf), np.full(len(graph_dims), -np.inf)])
    for data in (d for split in splits for d in split):
        for k, (a, b) in enumerate(graph_rng):
            graph_mm[0, k] = min(graph_mm[0, k], float(data.y[a:b].min()))
            graph_mm[1, k] = max(graph_mm[1, k], float(data.y[a:b].max()))
        for k, (a, b) in enumerate(node_rng):
            node_mm[0, k] = min(node_mm[0, k], float(data.x[:, a:b].min()))
            node_mm[1, k] = max(node_mm[1, k], float(data.x[:, a:b].max()))

    def scale(t, lo, hi):
        return (t - lo) / (hi - lo) if hi != lo else torch.zeros_like(t)

    for data in (d for split in splits for d in split):
        for k, (a, b) in enumerate(graph_rng):
            data.y[a:b] = scale(data.y[a:b], graph_mm[0, k], graph_mm[1, k])
        for k, (a, b) in enumerate(node_rng):
            data.x[:, a:b] = scale(data.x[:, a:b], node_mm[0, k], node_mm[1, k])
    return node_mm, graph_mm


def transform_raw_data_to_serialized(ds_config):
    """Parse the raw folders of `Dataset.path`, normalise, pickle one file per split."""
    node, graph = ds_config["node_features"], ds_config["graph_features"]
    out_dir = os.path.join(os.environ["SERIALIZED_DATA_PATH"], "serialized_dataset")
    os.makedirs(out_dir, exist_ok=True)
    names, splits = [], []
    for split, folder in ds_config["path"].items():
        folder = folder if os.path.isabs(folder) else os.path.join(os.getcwd(), folder)
        if not os.path.exists(folder):
            raise ValueError("Folder not found: ", folder)
        files = sorted(f for f in os.listdir(folder) if f != ".DS_Store" and os.path.isfile(os.path.join(folder, f)))
        assert files, "No data files provided in {}!".format(folder)
        samples = [_parse_lsms_file(os.path.join(folder, f), node["dim"], node["column_index"],
                                    graph["dim"], graph["column_index"]) for f in files]
        for k, nm in enumerate(graph["name"]):
            if "_scaled_num_nodes" in nm:
                for d in samples:
                    d.y[k] = d.y[k] / d.num_nodes
        splits.append(samples)
        names.append(ds_config["name"] + (".pkl" if split == "total" else f"_{split}.pkl"))
    node_mm, graph_mm = _minmax_normalize(splits, node["dim"], graph["dim"])
    for fname, samples in zip(names, splits):
        with open(os.path.join(out_dir, fname), "wb") as fh:
            pickle.dump(node_mm, fh)
            pickle.dump(graph_mm, fh)
            pickle.dump(samples, fh)


# --------------------------------------------------------- graph construction (GPU) ---
def _laplacian_pe(edge_index, num_nodes, k):
    """k lowest non-trivial eigenvectors of the symmetric normalised Laplacian (PyG
    AddLaplacianEigenvectorPE; signs fixed so the first non-zero entry is positive -- PyG draws
    random signs, which are not reproducible).  Zero-padded when the graph has < k+1 nodes."""
    pe = torch.zeros(num_nodes, k)
    if k == 0 or num_nodes < 2:
        return pe
    A = torch.zeros(num_nodes, num_nodes, dtype=torch.float64)
    A[edge_index[0], edge_index[1]] = 1.0
    A = torch.maximum(A, A.t())
    deg = A.sum(1).clamp(min=1e-12)
    L = torch.eye(num_nodes, dtype=torch.float64) - A / torch.sqrt(deg[:, None] * deg[None, :])
    vals, vecs = torch.linalg.eigh(L)
    vecs = vecs[:, 1:k + 1]
    for j in range(vecs.size(1)):
        nz = torch.nonzero(vecs[:, j].abs() > 1e-9)
        if nz.numel() and vecs[nz[0, 0], j] < 0:
            vecs[:, j] = -vecs[:, j]
    pe[:, :vecs.size(1)] = vecs.to(torch.float32)
    return pe


def _laplacian_pe_batched(samples, k, device):
    """`_laplacian_pe` for every sample, on `device`, one batched eigendecomposition per graph size
    (datasets of fixed-size structures: one call for the whole split instead of one dense fp64 eigh
    per sample on the host; chunked to bound memory)."""
    by_size = {}
    for i, d in enumerate(samples):
        by_size.setdefault(d.num_nodes, []).append(i)
    for n, idxs in by_size.items():
        if k == 0 or n < 2:
            for i in idxs:
                samples[i].pe = torch.zeros(n, k)
            continue
        chunk = max(1, min(len(idxs), (1 << 27) // (n * n)))          # <= 1 GiB of fp64 matrices
        for c0 in range(0, len(idxs), chunk):
            part = idxs[c0:c0 + chunk]
            A = torch.zeros((len(part), n, n), dtype=torch.float64, device=device)
            for b, i in enumerate(part):
                ei = samples[i].edge_index.to(device)
                A[b, ei[0], ei[1]] = 1.0
            A = torch.maximum(A, A.transpose(1, 2))
            deg = A.sum(2).clamp(min=1e-12)
            L = torch.eye(n, dtype=torch.float64, device=device) - A / torch.sqrt(deg[:, :, None] * deg[:, None, :])
            vecs = torch.linalg.eigh(L).eigenvectors[:, :, 1:k + 1]                  # [b, n, <=k]
            # sign: first entry with |v| > 1e-9 positive
            big = vecs.abs() > 1e-9
            first = torch.argmax(big.to(torch.int8), dim=1, keepdim=True)           # [b, 1, k']
            lead = torch.gather(vecs, 1, first)
            sign = torch.where((lead < 0) & big.any(dim=1, keepdim=True), -1.0, 1.0)
            vecs = (vecs * sign).to(torch.float32).cpu()
            for b, i in enumerate(part):
                pe = torch.zeros(n, k)
                pe[:, :vecs.size(2)] = vecs[b]
                samples[i].pe = pe


def build_graphs(samples, arch, device):
    """Edges, edge lengths (normalised by the global maximum) and encodings for a list of samples,
    with ONE batched radius-graph call on `device` (reference: one call per sample on the CPU)."""
    from ..ops import radius_graph, radius_graph_pbc

    if not samples:
        return samples
    sizes = [d.num_nodes for d in samples]
    ptr = torch.tensor([0] + list(np.cumsum(sizes)), dtype=torch.long)
    pos = torch.cat([d.pos for d in samples]).to(device)
    pbc = bool(arch.get("periodic_boundary_conditions"))
    cell = pbc_flags = None
    if pbc:
        cell = torch.stack([d.cell.reshape(3, 3) for d in samples])
        pbc_flags = torch.ones(len(samples), 3, dtype=torch.bool)
    if pbc:   # RadiusGraphPBC incl. cutoff expansion / artificial edges for unconnected nodes
        g = radius_graph_pbc(pos, ptr, arch["radius"], arch["max_neighbours"], loop=False, cell=cell, pbc=pbc_flags)
        edge_ptr = g["edge_ptr"]
    else:     # edges are grouped by destination, i.e. by sample
        g = radius_graph(pos, ptr, arch["radius"], arch["max_neighbours"], loop=False)
        edge_ptr = g["rowptr"].long().cpu()[ptr]
    src, dst = g["src"].long(), g["dst"].long()
    vec = pos[dst] - pos[src]
    if pbc:
        vec = vec + g["edge_shifts"]
    length = vec.norm(dim=1, keepdim=True)
    src, dst, length = src.cpu(), dst.cpu(), length.cpu()
    shifts = g["edge_shifts"].cpu() if pbc else None
    for k, d in enumerate(samples):
        a, b = int(edge_ptr[k]), int(edge_ptr[k + 1])
        d.edge_index = torch.stack([src[a:b], dst[a:b]]) - int(ptr[k])
        d.edge_attr = length[a:b].clone()
        if pbc:
            d.edge_shifts = shifts[a:b].clone()
    return samples


def finish_samples(samples, config, max_edge_length):
    """Edge normalisation, encodings, target / input selection (serialized_dataset_loader.py:153-205)."""
    arch = config["NeuralNetwork"]["Architecture"]
    var = config["NeuralNetwork"]["Variables_of_interest"]
    pe_dim = int(arch.get("pe_dim") or 0)
    node_dims = config["Dataset"]["node_features"]["dim"]
    graph_dims = config["Dataset"]["graph_features"]["dim"]
    if pe_dim > 0 and samples:
        device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
        _laplacian_pe_batched(samples, pe_dim, device)
    for d in samples:
        d.edge_attr = d.edge_attr / max_edge_length
        if pe_dim > 0:
            d.rel_pe = (d.pe[d.edge_index[0]] - d.pe[d.edge_index[1]]).abs()
        update_predicted_values(var["type"], var["output_index"], graph_dims, node_dims, d)
        d.x = d.x[:, list(var["input_node_features"])]
    return samples


def update_predicted_values(kinds, index, graph_feature_dim, node_feature_dim, data):
    """data.y <- concatenation of the selected graph / node targets; data.y_loc <- their offsets."""
    parts = []
    y_loc = torch.zeros(1, len(kinds) + 1, dtype=torch.int64)
    for k, (kind, idx) in enumerate(zip(kinds, index)):
        if kind == "graph":
            start = sum(graph_feature_dim[:idx])
            part = data.y[start:start + graph_feature_dim[idx]].reshape(-1, 1)
        elif kind == "node":
            start = sum(node_feature_dim[:idx])
            part = data.x[:, start:start + node_feature_dim[idx]].reshape(-1, 1)
        else:
            raise ValueError("Unknown output type", kind)
        parts.append(part)
        y_loc[0, k + 1] = y_loc[0, k] + part.numel()
    data.y, data.y_loc = torch.cat(parts, 0), y_loc


# ------------------------------------------------------------------------ loaders ---
class ListDataset(torch.utils.data.Dataset):
    def __init__(self, samples):
        self.dataset = list(samples)

    def __len__(self):
        return len(self.dataset)

    def __getitem__(self, idx):
        return self.dataset[idx]


def split_dataset(dataset, perc_train, stratify_splitting=False):
    """perc_train for training, the rest halved into validation and test (load_data.py:329-349)."""
    n = len(dataset)
    n_train = int(n * perc_train)
    n_val = int((n - n_train) / 2)
    return dataset[:n_train], dataset[n_train:n_train + n_val], dataset[n_train + n_val:]


def create_dataloaders(trainset, valset, testset, batch_size, train_sampler_shuffle=True,
                       test_sampler_shuffle=True):
    """Loaders of the three splits.  HYDRAGNN_DEVICE_COLLATE=1: every split is uploaded once into packed
    device arenas and the mini-batches are collated on the GPU (preprocess/packed_dataset.py) instead of
    on the host per batch."""
    size, rank = get_comm_size_and_rank()
    loaders = []
    if os.environ.get("HYDRAGNN_DEVICE_COLLATE", "0") == "1" and torch.cuda.is_available():
        from .packed_dataset import PackedDataset
        device = get_device()
        for samples, shuffle in ((trainset, train_sampler_shuffle), (valset, test_sampler_shuffle),
                                 (testset, test_sampler_shuffle)):
            samples = samples.dataset if isinstance(samples, ListDataset) else list(samples)
            ds = PackedDataset(samples, device)
            sampler = None
            if size > 1:
                sampler = torch.utils.data.distributed.DistributedSampler(ds, num_replicas=size, rank=rank,
                                                                          shuffle=shuffle)
            loaders.append(torch.utils.data.DataLoader(ds, batch_size=batch_size, shuffle=(sampler is None and shuffle),
                                                       sampler=sampler, collate_fn=ds.collate))
        return tuple(loaders)
    for samples, shuffle in ((trainset, train_sampler_shuffle), (valset, test_sampler_shuffle),
                             (testset, test_sampler_shuffle)):
        ds = samples if isinstance(samples, torch.utils.data.Dataset) else ListDataset(samples)
        sampler = None
        if size > 1:
            sampler = torch.utils.data.distributed.DistributedSampler(ds, num_replicas=size, rank=rank,
                                                                      shuffle=shuffle)
        loaders.append(torch.utils.data.DataLoader(
            ds, batch_size=batch_size, shuffle=(sampler is None and shuffle), sampler=sampler,
            collate_fn=Batch.from_data_list, pin_memory=torch.cuda.is_available()))
    return tuple(loaders)


def _load_pickle(path):
    with open(path, "rb") as fh:
        pickle.load(fh)
        pickle.load(fh)
        return pickle.load(fh)


def dataset_loading_and_splitting(config: dict):
    ds_config = config["Dataset"]
    arch = config["NeuralNetwork"]["Architecture"]
    training = config["NeuralNetwork"]["Training"]
    size, rank = get_comm_size_and_rank()
    paths = list(ds_config["path"].values())
    if not paths[0].endswith(".pkl"):
        if rank == 0:
            transform_raw_data_to_serialized(ds_config)
        if size > 1:
            torch.distributed.barrier()
        base = os.path.join(os.environ["SERIALIZED_DATA_PATH"], "serialized_dataset")
        paths = {k: os.path.join(base, ds_config["name"] + (".pkl" if k == "total" else f"_{k}.pkl"))
                 for k in ds_config["path"]}
    else:
        paths = dict(ds_config["path"])

    device = get_device()
    if device.type != "cuda":
        raise RuntimeError("graph construction runs on the GPU (libhgk.so); no CPU fallback")
    splits = {k: build_graphs(_load_pickle(p), arch, device) for k, p in paths.items()}
    max_len = max(float(d.edge_attr.max()) for samples in splits.values() for d in samples if d.edge_attr.numel())
    for k in splits:
        splits[k] = finish_samples(splits[k], config, max_len)
    if "total" in splits:
        samples = splits["total"]
        random.Random(0).shuffle(samples)
        trainset, valset, testset = split_dataset(samples, training["perc_train"],
                                                  ds_config.get("compositional_stratified_splitting", False))
    else:
        trainset, valset, testset = splits["train"], splits["validate"], splits["test"]
    return create_dataloaders(trainset, valset, testset, batch_size=training["batch_size"])
