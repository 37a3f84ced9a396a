"""Data-parallel plumbing: one process per GPU, NCCL over NVLink (gloo on CPU).

Same entry points as the reference (hydragnn/utils/distributed/distributed.py:106-198,
232-351): `setup_ddp`, `get_comm_size_and_rank`, `get_device(_name)`,
`get_distributed_model` (returns a wrapper exposing `.module`, which the train loop
dereferences, train_validate_test.py:94-96,523).

The gradient exchange is not torch DDP.  The models on this path hold 0.1-4 MB of
parameters, so the all-reduce is latency bound: all gradients live in ONE flat fp32
buffer (each `p.grad` is a view into it), cut into a few contiguous chunks in reverse
parameter order; a chunk's all-reduce is enqueued on a side stream the moment the last
gradient of the chunk has been accumulated, so it runs under the rest of backward.
`finish()` (called before the optimizer step) makes the compute stream wait for them.
Graphs are sharded by mini-batch per rank (DistributedSampler in the reference,
preprocess/load_data.py:251-274); there is no other data-path collective.
"""
import os

import torch
import torch.distributed as dist


def get_comm_size_and_rank():
    if dist.is_available() and dist.is_initialized():
        return int(dist.get_world_size()), int(dist.get_rank())
    return 1, 0


def get_local_rank():
    for key in ("LOCAL_RANK", "OMPI_COMM_WORLD_LOCAL_RANK", "SLURM_LOCALID"):
        if os.getenv(key):
            return int(os.environ[key])
    return 0


def _env_world():
    for size_key, rank_key in (("WORLD_SIZE", "RANK"), ("OMPI_COMM_WORLD_SIZE", "OMPI_COMM_WORLD_RANK"),
                               ("SLURM_NPROCS", "SLURM_PROCID")):
        if os.getenv(size_key) and os.getenv(rank_key):
            return int(os.environ[size_key]), int(os.environ[rank_key])
    return 1, 0


def setup_ddp(use_deepspeed=False):
    """env:// rendezvous; backend nccl when CUDA is present, else gloo (override: HYDRAGNN_BACKEND)."""
    if use_deepspeed:
        raise NotImplementedError("deepspeed is not part of the B200 path")
    world_size, rank = _env_world()
    if dist.is_initialized():
        return get_comm_size_and_rank()
    backend = os.getenv("HYDRAGNN_BACKEND") or ("nccl" if torch.cuda.is_available() else "gloo")
    os.environ.setdefault("MASTER_ADDR", os.getenv("HYDRAGNN_MASTER_ADDR", "127.0.0.1"))
    os.environ.setdefault("MASTER_PORT", os.getenv("HYDRAGNN_MASTER_PORT", "8889"))
    os.environ["WORLD_SIZE"], os.environ["RANK"] = str(world_size), str(rank)
    if backend == "nccl":
        torch.cuda.set_device(get_local_rank() % torch.cuda.device_count())
    if world_size > 1 or os.getenv("HYDRAGNN_FORCE_DIST"):
        kwargs = {}
        if backend == "nccl":
            kwargs["device_id"] = torch.device("cuda", torch.cuda.current_device())
        dist.init_process_group(backend=backend, init_method="env://", **kwargs)
    return world_size, rank


def get_device_name(use_gpu=True, rank_per_model=1, verbosity_level=0, no_prefix=False):
    if use_gpu and torch.cuda.is_available():
        idx = get_local_rank() % torch.cuda.device_count()
        return str(idx) if no_prefix else f"cuda:{idx}"
    return "cpu"


def get_device(use_gpu=True, rank_per_model=1, verbosity_level=0):
    return torch.device(get_device_name(use_gpu, rank_per_model, verbosity_level))


def is_model_distributed(model):
    return isinstance(model, (DataParallelModel, torch.nn.parallel.DistributedDataParallel))


def nsplit(a, n):
    k, m = divmod(len(a), n)
    return (a[i * k + min(i, m):(i + 1) * k + min(i + 1, m)] for i in range(n))


def comm_reduce(x, op):
    """All-reduce a host scalar / array with the process group (the reference uses mpi4py here)."""
    size, _ = get_comm_size_and_rank()
    t = torch.as_tensor(x, dtype=torch.float64)
    if size > 1:
        dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else "cpu"
        t = t.to(dev)
        dist.all_reduce(t, op=op)
        t = t.cpu()
    return t.numpy() if t.dim() else t.item()


class GradientAverager:
    """Gradient mean across ranks through one flat fp32 buffer, overlapped with backward.

    Gradients are NOT accumulated into the buffer by autograd (that costs one add kernel per
    parameter per step): `zero_grad` sets them to None, so backward hands each parameter its
    gradient tensor as is.  With more than one rank, a post-accumulate hook counts the parameters
    of a chunk (reverse registration order = the order backward produces them); when a chunk is
    complete its gradients are packed into the chunk's slice of the flat buffer with one `cat`,
    the parameters' `.grad` are re-pointed at views of that slice, and the slice is all-reduced
    asynchronously on a side stream while backward continues.  `finish()` waits, divides by the
    world size, and leaves `.grad` = averaged views for the optimizer.  With one rank nothing is
    packed at all.

    Collectives are matched across ranks by issue order, so the order must not depend on which
    parameters a rank's local batch happened to reach (a multibranch model runs only the heads
    present in its batch): chunk k is launched only after chunks 0..k-1, and `finish()` flushes
    what is left in index order.  A chunk with an unused parameter therefore waits for `finish()`
    on that rank and holds the later chunks back with it -- same sequence of all-reduces on
    every rank, whatever each one's autograd graph looked like."""

    def __init__(self, params, world_size=None, num_chunks=4, process_group=None):
        self.params = [p for p in params if p.requires_grad]
        self.world_size = world_size if world_size is not None else get_comm_size_and_rank()[0]
        self.group = process_group
        self.enabled = self.world_size > 1
        total = sum(p.numel() for p in self.params)
        dev = self.params[0].device if self.params else torch.device("cpu")
        self.flat = torch.zeros(total if self.enabled else 0, dtype=torch.float32, device=dev)
        # reverse order: the last layers' gradients are produced first by backward
        order = list(reversed(self.params))
        per_chunk = max(1, -(-total // max(1, num_chunks)))
        self.chunks, self._chunk_of = [], {}
        off, start, members = 0, 0, []
        for p in order:
            members.append(p)
            off += p.numel()
            if off - start >= per_chunk:
                self._close_chunk(start, off, members)
                start, members = off, []
        if members:
            self._close_chunk(start, off, members)
        self._pending = [len(c["members"]) for c in self.chunks]
        self._works = []
        self._next = 0              # chunks [0, _next) have been launched this step
        # NCCL averages inside the collective (ReduceOp.AVG): no separate scaling pass over the buffer after
        # the last chunk has arrived; gloo (CPU tests) has no AVG: SUM, then one multiply in finish()
        self._avg_in_collective = bool(self.enabled and dev.type == "cuda" and dist.is_initialized()
                                       and dist.get_backend(self.group) == "nccl")
        self.stream = torch.cuda.Stream(device=dev) if (dev.type == "cuda" and self.enabled) else None
        for p in self.params:
            p.grad = None
        if self.enabled:
            for p in self.params:
                p.register_post_accumulate_grad_hook(self._on_grad_ready)

    def _close_chunk(self, start, stop, members):
        idx = len(self.chunks)
        views, off = [], start
        if self.enabled:
            for p in members:
                views.append(self.flat[off:off + p.numel()].view_as(p))
                off += p.numel()
        self.chunks.append({"view": self.flat[start:stop] if self.enabled else None, "members": list(members),
                            "views": views})
        for p in members:
            self._chunk_of[p] = idx

    def zero_grad(self):
        for p in self.params:
            p.grad = None
        self._pending = [len(c["members"]) for c in self.chunks]
        self._works = []
        self._next = 0

    def _launch(self, idx):
        chunk = self.chunks[idx]
        # pack (parameters backward never reached contribute zeros), then hand .grad the views
        parts = [(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1) for p in chunk["members"]]
        torch.cat(parts, out=chunk["view"])
        for p, v in zip(chunk["members"], chunk["views"]):
            p.grad = v
        op = dist.ReduceOp.AVG if self._avg_in_collective else dist.ReduceOp.SUM
        if self.stream is not None:
            self.stream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self.stream):
                work = dist.all_reduce(chunk["view"], op=op, group=self.group, async_op=True)
        else:
            work = dist.all_reduce(chunk["view"], op=op, group=self.group, async_op=True)
        self._works.append(work)

    suspended = False      # set by train/step_graph.py, which reduces one flat buffer itself

    def _on_grad_ready(self, p):
        if self.suspended:
            return
        idx = self._chunk_of[p]
        self._pending[idx] -= 1
        # rank-independent order: only the chunk at the cursor may go, then any completed ones behind it
        while self._next < len(self.chunks) and self._pending[self._next] == 0:
            self._launch(self._next)
            self._next += 1

    def finish(self):
        """Wait for the chunk all-reduces, launch any chunk backward never completed (unused
        parameters), scale by 1/world_size.  Call after backward, before optimizer.step()."""
        if not self.enabled:
            return
        while self._next < len(self.chunks):
            self._launch(self._next)
            self._pending[self._next] = 0
            self._next += 1
        for w in self._works:
            w.wait()
        if self.stream is not None:
            torch.cuda.current_stream().wait_stream(self.stream)
        if not self._avg_in_collective:
            self.flat.mul_(1.0 / self.world_size)
        self._works = []


class DataParallelModel(torch.nn.Module):
    """`.module` + forward pass-through, with the GradientAverager attached."""

    def __init__(self, module, num_chunks=4):
        super().__init__()
        self.module = module
        size, _ = get_comm_size_and_rank()
        if size > 1:
            # same initial weights everywhere (torch DDP broadcasts rank 0's at construction too)
            for t in list(module.parameters()) + list(module.buffers()):
                dist.broadcast(t.data, src=0)
        self.averager = GradientAverager(list(module.parameters()), size, num_chunks)

    def forward(self, *args, **kwargs):
        return self.module(*args, **kwargs)

    def zero_grad(self, set_to_none=False):
        self.averager.zero_grad()

    def finish_gradients(self):
        self.averager.finish()


def get_distributed_model(model, verbosity=0, sync_batch_norm=False, find_unused_parameters=False):
    size, _ = get_comm_size_and_rank()
    if sync_batch_norm and size > 1:
        model = torch.nn.SyncBatchNorm.convert_sync_batchnorm(model)
    return DataParallelModel(model)
