"""CPU, gloo, world_size 2: the data-parallel plumbing (hydragnn.utils.distributed) -- same initial
weights on every rank, gradients averaged across ranks through the flat-buffer GradientAverager,
identical parameters after the optimizer step."""
import copy
import os
import socket
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, os.path.join(ROOT, "hydragnn-gps-support_b200"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank), HYDRAGNN_BACKEND="gloo")
    from hydragnn.utils.distributed import get_distributed_model, setup_ddp

    size, r = setup_ddp()
    assert (size, r) == (world, rank)
    torch.manual_seed(100 + rank)                      # different init per rank on purpose
    net = torch.nn.Sequential(torch.nn.Linear(6, 16), torch.nn.ReLU(), torch.nn.Linear(16, 3), torch.nn.Linear(3, 1))
    model = get_distributed_model(net)                 # broadcasts rank 0's weights
    opt = torch.optim.SGD(net.parameters(), lr=0.1)
    torch.manual_seed(7 + rank)
    x, y = torch.randn(5 + rank, 6), torch.randn(5 + rank, 1)    # each rank its own shard
    # this rank's own gradient, from an unwrapped copy: the averager starts reducing chunks of the flat
    # buffer while backward is still running, so p.grad of the wrapped net is not "local" any more
    twin = copy.deepcopy(net)
    torch.nn.functional.mse_loss(twin(x), y).backward()
    local = [p.grad.clone() for p in twin.parameters()]
    model.zero_grad()
    loss = torch.nn.functional.mse_loss(model(x), y)
    loss.backward()
    model.finish_gradients()
    averaged = [p.grad.clone() for p in net.parameters()]
    opt.step()
    torch.save({"local": local, "avg": averaged, "params": [p.detach().clone() for p in net.parameters()]},
               os.path.join(out_dir, f"rank{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_gradient_averager_gloo_world2(tmp_path):
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    r0 = torch.load(tmp_path / "rank0.pt")
    r1 = torch.load(tmp_path / "rank1.pt")
    for g0, g1, a0, a1 in zip(r0["local"], r1["local"], r0["avg"], r1["avg"]):
        torch.testing.assert_close(a0, (g0 + g1) / 2, rtol=1e-6, atol=1e-7)
        assert torch.equal(a0, a1)
    for p0, p1 in zip(r0["params"], r1["params"]):
        assert torch.equal(p0, p1)                      # replicas stay in lock step


class _TwoHeads(torch.nn.Module):
    """Trunk + two heads; a rank runs only the heads named in `use` (Base._forward runs only the
    branch heads present in the local batch)."""

    def __init__(self):
        super().__init__()
        self.trunk = torch.nn.Linear(6, 8)
        self.head_a = torch.nn.Linear(8, 1)
        self.head_b = torch.nn.Linear(8, 1)

    def forward(self, x, use):
        h = torch.relu(self.trunk(x))
        out = 0.0
        if "a" in use:
            out = out + self.head_a(h)
        if "b" in use:
            out = out + self.head_b(h)
        return out


def _worker_unused_head(rank, world, port, out_dir):
    sys.path.insert(0, os.path.join(ROOT, "hydragnn-gps-support_b200"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank), HYDRAGNN_BACKEND="gloo")
    from hydragnn.utils.distributed import DataParallelModel, setup_ddp

    setup_ddp()
    torch.manual_seed(3)
    net = _TwoHeads()
    # heads and trunk in separate chunks: with completion-order launches rank 0 (head_b unused) and
    # rank 1 (all used) would issue their all-reduces in different orders -- mismatched sizes, a hang
    model = DataParallelModel(net, num_chunks=64)
    chunk_of = model.averager._chunk_of
    assert len({chunk_of[net.head_a.weight], chunk_of[net.head_b.weight], chunk_of[net.trunk.weight]}) == 3
    torch.manual_seed(11 + rank)
    x, y = torch.randn(4, 6), torch.randn(4, 1)
    use = ("a",) if rank == 0 else ("a", "b")
    twin = copy.deepcopy(net)
    torch.nn.functional.mse_loss(twin(x, use), y).backward()
    local = {n: (p.grad.clone() if p.grad is not None else torch.zeros_like(p)) for n, p in twin.named_parameters()}
    model.zero_grad()
    torch.nn.functional.mse_loss(model(x, use), y).backward()
    model.finish_gradients()
    torch.save({"local": local, "avg": {n: p.grad.clone() for n, p in net.named_parameters()}},
               os.path.join(out_dir, f"rank{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_gradient_averager_rank_with_unused_head(tmp_path):
    """A rank whose batch leaves a head unused issues the same sequence of collectives."""
    world, port = 2, _free_port()
    mp.spawn(_worker_unused_head, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    r0 = torch.load(tmp_path / "rank0.pt")
    r1 = torch.load(tmp_path / "rank1.pt")
    for n in r0["local"]:
        torch.testing.assert_close(r0["avg"][n], (r0["local"][n] + r1["local"][n]) / 2, rtol=1e-6, atol=1e-7)
        assert torch.equal(r0["avg"][n], r1["avg"][n]), n
