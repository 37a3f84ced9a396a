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
