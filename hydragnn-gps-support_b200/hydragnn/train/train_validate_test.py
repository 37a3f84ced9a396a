"""Training loop on the B200 path.

Control flow and names follow the reference's hydragnn/train/train_validate_test.py
(`train_validate_test`:52, `train`:451, `validate`:570, `test`:620, `get_head_indices`:316):
per batch -- zero grads, head indices, move to device, forward, loss (or the energy+force loss
with create_graph), backward, optimizer step; per epoch -- rank-averaged losses.
Differences that matter for throughput: head indices are computed without Python loops over
samples or device syncs (the batch carries `num_graphs` / `y_loc` from the collate), and the
gradient all-reduce is the flat-buffer GradientAverager overlapped with backward.
"""
import os

import torch
import torch.distributed as dist

from ..utils.distributed import get_comm_size_and_rank


def _core(model):
    return model.module if hasattr(model, "module") else model


def _num_graphs(data):
    n = getattr(data, "num_graphs", None)
    if n is None:
        n = int(data.batch.max()) + 1
    return int(n)


def get_head_indices(model, data):
    """Index of every head's targets inside the flat `data.y` of a mini-batch.

    `data.y` concatenates, per graph, the targets of all heads (graph heads: head_dim values;
    node heads: num_nodes * head_dim values, located by the cumulative `y_loc` [B, heads+1])."""
    core = _core(model)
    dev = data.y.device
    if core.num_heads == 1:
        return [torch.arange(data.y.shape[0], device=dev)]
    B = _num_graphs(data)
    if all(kind == "graph" for kind in core.head_type):
        dims = list(core.head_dims)
        stride = sum(dims)
        base = torch.arange(B, device=dev).view(-1, 1) * stride
        out, off = [], 0
        for d in dims:
            out.append((base + off + torch.arange(d, device=dev).view(1, -1)).reshape(-1))
            off += d
        return out
    y_loc = data.y_loc.to(dev).long().view(B, -1)
    total = y_loc[:, -1]
    start_of_graph = torch.cumsum(total, 0) - total
    out = []
    for h in range(core.num_heads):
        lo = start_of_graph + y_loc[:, h]
        length = y_loc[:, h + 1] - y_loc[:, h]
        # ranges lo[g] .. lo[g]+length[g], concatenated over graphs, without a Python loop
        seg = torch.repeat_interleave(torch.arange(B, device=dev), length)
        pos_in_seg = torch.arange(seg.numel(), device=dev) - (torch.cumsum(length, 0) - length)[seg]
        out.append(lo[seg] + pos_in_seg)
    return out


def train_step(model, optimizer, data, compute_grad_energy=False, head_index=None):
    """One optimisation step on a batch that already lives on the model's device.
    Returns the (detached) loss tensor and the per-task losses; no host sync."""
    core = _core(model)
    if hasattr(model, "averager"):
        model.zero_grad()
    else:
        optimizer.zero_grad(set_to_none=False)
    if compute_grad_energy:
        data.pos.requires_grad_(True)
        pred = model(data)
        loss, tasks_loss = core.energy_force_loss(pred, data)
    else:
        if head_index is None:
            head_index = get_head_indices(model, data)
        pred = model(data)
        loss, tasks_loss = core.loss(pred, data.y, head_index)
    loss.backward()
    if hasattr(model, "finish_gradients"):
        model.finish_gradients()
    optimizer.step()
    return loss.detach(), tasks_loss


def _step_graph_for(model, opt):
    """One BucketedStepGraph per (model, optimizer), kept on the model object.  The optimizer is switched
    to its capturable form (device-side step counter) before its first step."""
    from .step_graph import BucketedStepGraph

    sg = getattr(model, "_hgk_step_graph", None)
    if sg is None or sg.optimizer is not opt:
        if not opt.state:
            for group in opt.param_groups:
                if "capturable" in group:
                    group["capturable"] = True
        sg = BucketedStepGraph(model, opt, node_bucket=int(os.getenv("HYDRAGNN_STEP_GRAPH_NODES", "64")),
                               edge_bucket=int(os.getenv("HYDRAGNN_STEP_GRAPH_EDGES", "1024")),
                               max_entries=int(os.getenv("HYDRAGNN_STEP_GRAPH_MAX", "64")))
        try:
            model._hgk_step_graph = sg
        except Exception:
            pass
    return sg


def _reduce_mean(values):
    size, _ = get_comm_size_and_rank()
    if size > 1:
        dist.all_reduce(values, op=dist.ReduceOp.SUM)
        values = values / size
    return values


def train(loader, model, opt, verbosity=0, profiler=None, use_deepspeed=False, compute_grad_energy=False):
    core = _core(model)
    device = next(core.parameters()).device
    model.train()
    total = torch.zeros(1, device=device)
    tasks = torch.zeros(core.num_heads, device=device)
    seen = 0
    max_batches = int(os.getenv("HYDRAGNN_MAX_NUM_BATCH", "0")) or None
    step_graph = _step_graph_for(model, opt) if os.getenv("HYDRAGNN_STEP_GRAPH", "0") == "1" else None
    for ibatch, data in enumerate(loader):
        if max_batches is not None and ibatch >= max_batches:
            break
        data = data.to(device)
        if step_graph is not None:      # whole steps replayed as CUDA graphs over size buckets (step_graph.py)
            loss, tasks_loss = step_graph.step(data, compute_grad_energy)
        else:
            loss, tasks_loss = train_step(model, opt, data, compute_grad_energy)
        n = _num_graphs(data)
        total += loss * n
        for k, t in enumerate(tasks_loss):
            tasks[k] += t.detach() * n
        seen += n
        if profiler is not None:
            profiler.step()
    seen = max(seen, 1)
    return _reduce_mean(total / seen).item(), _reduce_mean(tasks / seen)


@torch.no_grad()
def _evaluate(loader, model, compute_grad_energy, collect):
    core = _core(model)
    device = next(core.parameters()).device
    model.eval()
    total = torch.zeros(1, device=device)
    tasks = torch.zeros(core.num_heads, device=device)
    seen = 0
    true_values = [[] for _ in range(core.num_heads)]
    pred_values = [[] for _ in range(core.num_heads)]
    for data in loader:
        data = data.to(device)
        if compute_grad_energy:
            with torch.enable_grad():
                data.pos.requires_grad_(True)
                pred = model(data)
                loss, tasks_loss = core.energy_force_loss(pred, data)
            loss = loss.detach()
        else:
            head_index = get_head_indices(model, data)
            pred = model(data)
            loss, tasks_loss = core.loss(pred, data.y, head_index)
            if collect:
                pred_heads = pred[0] if core.var_output else pred
                for k in range(core.num_heads):
                    true_values[k].append(data.y[head_index[k]].reshape(-1, 1))
                    pred_values[k].append(pred_heads[k].reshape(-1, 1))
        n = _num_graphs(data)
        total += loss * n
        for k, t in enumerate(tasks_loss):
            tasks[k] += t.detach() * n
        seen += n
    seen = max(seen, 1)
    error = _reduce_mean(total / seen).item()
    tasks = _reduce_mean(tasks / seen)
    if collect:
        true_values = [torch.cat(v) if v else torch.zeros(0, 1, device=device) for v in true_values]
        pred_values = [torch.cat(v) if v else torch.zeros(0, 1, device=device) for v in pred_values]
        true_values = [gather_tensor_ranks(v) for v in true_values]
        pred_values = [gather_tensor_ranks(v) for v in pred_values]
    return error, tasks, true_values, pred_values


def validate(loader, model, verbosity=0, reduce_ranks=True, compute_grad_energy=False):
    error, tasks, _, _ = _evaluate(loader, model, compute_grad_energy, collect=False)
    return error, tasks


def test(loader, model, verbosity=0, reduce_ranks=True, return_samples=True, compute_grad_energy=False):
    return _evaluate(loader, model, compute_grad_energy, collect=return_samples)


def gather_tensor_ranks(head_values):
    """Concatenate a per-rank [n_local, d] tensor over ranks (pad to the longest, gather, trim)."""
    size, _ = get_comm_size_and_rank()
    if size == 1:
        return head_values
    n_local = torch.tensor([head_values.shape[0]], device=head_values.device)
    counts = [torch.zeros_like(n_local) for _ in range(size)]
    dist.all_gather(counts, n_local)
    n_max = int(max(c.item() for c in counts))
    padded = torch.zeros((n_max,) + tuple(head_values.shape[1:]), dtype=head_values.dtype,
                         device=head_values.device)
    padded[:head_values.shape[0]] = head_values
    parts = [torch.zeros_like(padded) for _ in range(size)]
    dist.all_gather(parts, padded)
    return torch.cat([p[:int(c.item())] for p, c in zip(parts, counts)], dim=0)


def train_validate_test(model, optimizer, train_loader, val_loader, test_loader, writer, scheduler,
                        config, model_with_config_name, verbosity=0, plot_init_solution=True,
                        plot_hist_solution=False, create_plots=False, use_deepspeed=False,
                        compute_grad_energy=False):
    """Epoch loop: train, validate, test, LR schedule, checkpoint / early stopping."""
    from ..utils.model import Checkpoint, EarlyStopping

    training = config["Training"]
    num_epoch = training["num_epoch"]
    early = EarlyStopping(patience=training.get("patience", 10)) if training.get("EarlyStopping") else None
    ckpt = None
    if training.get("Checkpoint"):
        ckpt = Checkpoint(name=model_with_config_name, warmup=training.get("checkpoint_warmup", 0))
    history = {"train": [], "val": [], "test": []}
    run_valtest = bool(int(os.getenv("HYDRAGNN_VALTEST", "1")))
    for epoch in range(training.get("epoch_start", 0), num_epoch):
        for loader in (train_loader, val_loader, test_loader):
            sampler = getattr(loader, "sampler", None)
            if hasattr(sampler, "set_epoch"):
                sampler.set_epoch(epoch)
        train_loss, _ = train(train_loader, model, optimizer, verbosity, compute_grad_energy=compute_grad_energy)
        history["train"].append(train_loss)
        if not run_valtest:
            continue
        val_loss, _ = validate(val_loader, model, verbosity, compute_grad_energy=compute_grad_energy)
        test_loss, _, _, _ = test(test_loader, model, verbosity, return_samples=False,
                                  compute_grad_energy=compute_grad_energy)
        history["val"].append(val_loss)
        history["test"].append(test_loss)
        if scheduler is not None:
            scheduler.step(val_loss)
        if writer is not None:
            writer.add_scalar("train error", train_loss, epoch)
            writer.add_scalar("validate error", val_loss, epoch)
            writer.add_scalar("test error", test_loss, epoch)
        if verbosity >= 2 and get_comm_size_and_rank()[1] == 0:
            print(f"Epoch: {epoch:02d}, Train Loss: {train_loss:.8f}, Val Loss: {val_loss:.8f}, "
                  f"Test Loss: {test_loss:.8f}")
        if ckpt is not None and ckpt(model, optimizer, val_loss):
            pass
        if early is not None and early(val_loss):
            break
    return history
