"""`-m gpu`: device-side collate (hydragnn/preprocess/packed_dataset.py on hgk_collate_rows / hgk_collate_edges)
against the host collate it replaces (Batch.from_data_list, the stand-in for torch_geometric's, reference
hydragnn/preprocess/load_data.py:225-326): every field of the mini-batch bit for bit -- index / byte work."""
import copy
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


def _samples(n_samples, seed, with_pbc=True):
    from hydragnn.graph import Data

    gen = torch.Generator().manual_seed(seed)
    out = []
    for i in range(n_samples):
        n = int(torch.randint(1, 12, (1,), generator=gen))
        e = int(torch.randint(0, 4 * n + 1, (1,), generator=gen)) if i % 5 else 0      # some samples without edges
        d = Data(x=torch.randn(n, 3, generator=gen), pos=torch.randn(n, 3, generator=gen),
                 pe=torch.randn(n, 4, generator=gen), forces=torch.randn(n, 3, generator=gen),
                 edge_index=torch.randint(0, n, (2, e), generator=gen),
                 edge_attr=torch.rand(e, 1, generator=gen), rel_pe=torch.rand(e, 4, generator=gen),
                 edge_shifts=torch.randn(e, 3, generator=gen),
                 y=torch.randn(1 + 2 * n, 1, generator=gen),
                 y_loc=torch.tensor([[0, 1, 1 + n, 1 + 2 * n]], dtype=torch.int64),
                 energy=torch.randn(1, generator=gen), dataset_name=torch.tensor([[i % 3]], dtype=torch.int64))
        if with_pbc:
            d.cell = torch.randn(3, 3, generator=gen)
            d.pbc = torch.tensor([True, i % 2 == 0, False])
        out.append(d)
    return out


@pytest.mark.parametrize("indices", [[0], [3, 1, 4, 1, 5], list(range(40)), [39, 0, 20, 20, 5, 10]])
def test_device_collate_equals_host_collate(libhgk, indices):
    from hydragnn.graph import Batch
    from hydragnn.preprocess.packed_dataset import PackedDataset

    samples = _samples(40, 7)
    packed = PackedDataset(samples, "cuda:0")
    got = packed.collate(indices)
    want = Batch.from_data_list([samples[i] for i in indices]).to("cuda:0")
    assert got.num_graphs == want.num_graphs and got.graph_sizes == want.graph_sizes
    keys = [k for k in want.keys() if torch.is_tensor(getattr(want, k))]
    assert set(keys) <= set(got.keys()), set(keys) - set(got.keys())
    for k in keys:
        a, b = getattr(got, k), getattr(want, k)
        assert a.shape == b.shape and a.dtype == b.dtype, (k, a.shape, b.shape, a.dtype, b.dtype)
        assert torch.equal(a, b), k


def test_training_with_device_collate(libhgk, tmp_path, monkeypatch):
    """run_training / run_prediction on the CI toy with HYDRAGNN_DEVICE_COLLATE=1 reach the reference's gate."""
    import hydragnn
    import toy_data
    from test_run_training_gpu import CI_CONFIG

    monkeypatch.chdir(tmp_path)
    monkeypatch.setenv("SERIALIZED_DATA_PATH", str(tmp_path))
    monkeypatch.setenv("HYDRAGNN_DEVICE_COLLATE", "1")
    for split, count, seed in (("train", 350, 1), ("test", 75, 2), ("validate", 75, 3)):
        toy_data.generate(os.path.join(tmp_path, "dataset", f"unit_test_singlehead_{split}"), count, seed)
    config = copy.deepcopy(CI_CONFIG)
    config["NeuralNetwork"]["Architecture"]["mpnn_type"] = "SchNet"
    hydragnn.run_training(copy.deepcopy(config))
    error, task_errors, true_values, predicted = hydragnn.run_prediction(copy.deepcopy(config))
    assert true_values[0].shape[0] == 75
    assert torch.sqrt(torch.mean((true_values[0] - predicted[0]) ** 2)).item() < 0.20
