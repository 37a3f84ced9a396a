"""Whole training steps as CUDA graphs, keyed on bucketed batch sizes.

For small-molecule workloads (BASELINE configs 2 and 3: ~1000 nodes, 3-15 k edges per mini-batch) a
training step is 1-3 ms of kernels behind 5-13 ms of host work: a few hundred to a few thousand
launches and autograd nodes.  The reference pays the same price (it is eager PyTorch).  Here the
whole step -- forward, loss, backward, optimizer -- is captured once per SIZE BUCKET and replayed:

  * a mini-batch is padded to (N_b, E_b): N_b = N + >= 2 pad nodes rounded up to `node_bucket`,
    E_b = E rounded up to `edge_bucket`.  The pad nodes form one extra graph (id B) of their own; the
    pad edges connect pad nodes only.  Nothing real ever reads a pad row: message passing follows the
    edge list, pooling follows `batch`, attention follows `cu_seqlens` ([0, N, N_b]: the pad nodes are
    a segment of their own), and BatchNorm statistics of the GPS block count the first N rows only
    (`ops.node.valid_rows`, read from device memory by hgk_bn_stats / hgk_bn_bwd);
  * the loss runs over padded targets with the pad entries replaced by the (detached) predictions
    themselves, i.e. they contribute exactly zero to the loss and to every gradient, and the mean is
    rescaled by n_padded / n_real (a device scalar), so the value equals the unpadded loss;
  * tensors of one bucket live in static buffers; a step copies the batch in and replays the graph.
    Shapes are static per bucket, batch-dependent values (sizes, offsets, scale) are device tensors.

Data-parallel training: the captured part is forward + loss + backward, which packs the gradients into ONE
flat buffer shared by all buckets (the parameters' .grad are views of it); the gradient mean over the
ranks is one eager all-reduce of that buffer between the replay and a second, bucket-independent graph
that holds the optimizer step.  Warm-up and capture issue no collective, so ranks may capture at
different steps without falling out of step with each other.

A stream of random mini-batches of a molecular dataset hits a handful of buckets (sum of 64 molecule
sizes: +-2 %), each captured on first use.  Eligible: mse / mae losses, SchNet /
EGNN / PAINN stacks (their per-node normalisation, if any, is the GPS block's, which honours
`valid_rows`), with or without the energy + force loss (the double backward is captured like
everything else; not through a GPS layer); everything else runs the eager `train_step`.

Opt in with `HYDRAGNN_STEP_GRAPH=1` (train loop) or `BucketedStepGraph(model, optimizer)` directly.
"""
import torch

from .. import _hgk
from ..graph import Data
from ..ops import node as node_ops
from ..ops import radius_graph
from .train_validate_test import _core, get_head_indices, train_step

_NODE = ("x", "pos", "pe")
_EDGE = ("edge_attr", "rel_pe", "edge_shifts")


def _roundup(n, m):
    return (n + m - 1) // m * m


class _Entry:
    __slots__ = ("static", "graph", "loss", "tasks", "uses", "launches")


class BucketedStepGraph:
    def __init__(self, model, optimizer, node_bucket=64, edge_bucket=1024, max_entries=32, warmup=2):
        self.model, self.optimizer = model, optimizer
        self.core = _core(model)
        self.node_bucket, self.edge_bucket = int(node_bucket), int(edge_bucket)
        self.max_entries, self.warmup = max_entries, warmup
        self.entries = {}
        self.replays = self.captures = self.eager_steps = 0
        # The captured optimizer step must not bake the learning rate in as a constant (schedulers change
        # it): while capturing, the param groups carry the rate as a device tensor; before a replay the
        # tensor is refreshed from the (host) value the scheduler left in the group.
        self._lr_dev, self._lr_host = None, None
        self._flat = None              # flat fp32 gradient buffer; p.grad = views of it
        self._views = None
        self._opt_graph = None         # the optimizer step, captured once
        import torch.distributed as dist
        self._world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1

    # ----------------------------------------------------------------------------- eligibility ---
    def eligible(self, data, compute_grad_energy=False):
        core = self.core
        if not torch.cuda.is_available():
            return False
        if compute_grad_energy and (core.use_global_attn or core.num_heads != 1 or core.head_type[0] != "node"
                                    or getattr(data, "energy", None) is None or getattr(data, "forces", None) is None):
            return False            # the force loss through a GPS layer differentiates attention twice: eager
        if core.loss_function_type not in ("mse", "mae") or core.var_output or core.num_branches > 1:
            return False
        name = str(core)
        if name not in ("SCFStack", "EGCLStack", "PAINNStack"):
            return False
        if getattr(core, "equivariance", False) or getattr(core, "conv_checkpointing", False):
            return False
        if any(kind == "node" and self.core.config_heads["node"][0]["architecture"]["type"] != "mlp"
               for kind in core.head_type):
            return False
        if not all(pg.get("capturable", False) for pg in self.optimizer.param_groups):
            return False
        if core.use_global_attn:
            # BatchNorm of the GPS block must be the kernel-library one (it honours `valid_rows`)
            from ..globalAtt import gps
            if not (gps.NATIVE_TAIL and core.training and node_ops.supported_width(core.hidden_dim)):
                return False
        if not data.pos.is_cuda:
            return False
        return len(self.entries) < self.max_entries or self._key_of(data, compute_grad_energy) in self.entries

    # ------------------------------------------------------------------------------- padding ---
    def _own_graph(self):
        """Plain SchNet builds its radius graph inside the model (SCFStack.py:120-153)."""
        return str(self.core) == "SCFStack" and not self.core._edge_mode()

    def _edges_of(self, data):
        if not self._own_graph():
            return data.edge_index
        cached = getattr(data, "_step_graph_edges", None)
        if cached is None:
            ptr = data.ptr.to(device=data.pos.device, dtype=torch.int32)
            g = radius_graph(data.pos, ptr, self.core.radius, self.core.max_neighbours, loop=False)
            cached = data._step_graph_edges = g["edge_index"]
        return cached

    def _key_of(self, data, forces=False):
        N, B = data.pos.size(0), int(data.num_graphs)
        E = self._edges_of(data).size(1)
        return (_roundup(N + 2, self.node_bucket), _roundup(max(E, 1), self.edge_bucket), B, bool(forces))

    def _pad(self, data, key):
        """-> dict of padded tensors (+ masks, scale) for bucket `key`; all on the device."""
        N_b, E_b, B, forces = key
        dev = data.pos.device
        N = data.pos.size(0)
        ei = self._edges_of(data)
        E = ei.size(1)
        P, EP = N_b - N, E_b - E
        out = {}
        for k in _NODE:
            v = getattr(data, k, None)
            if v is None:
                continue
            pad = torch.zeros((P,) + tuple(v.shape[1:]), dtype=v.dtype, device=dev)
            if k == "pos":       # pad nodes in pairs one length unit apart, far from everything else
                idx = torch.arange(P, device=dev, dtype=v.dtype)
                pad[:, 0] = 1.0e3 + 3.0 * torch.div(idx, 2, rounding_mode="floor") + (idx % 2)
            out[k] = torch.cat([v, pad], 0)
        # pad edges: (N + 2m) <-> (N + 2m + 1), cycling over the pad pairs, receivers ascending
        pairs = max(P // 2, 1)
        m = torch.arange(EP, device=dev) % (2 * pairs)
        dst = torch.sort(N + m).values
        src = torch.where((dst - N) % 2 == 0, dst + 1, dst - 1)
        out["edge_index"] = torch.cat([ei, torch.stack([src, dst]).to(ei.dtype)], 1)
        for k in _EDGE:
            v = getattr(data, k, None)
            if v is not None and not self._own_graph():
                out[k] = torch.cat([v, torch.zeros((EP,) + tuple(v.shape[1:]), dtype=v.dtype, device=dev)], 0)
        out["batch"] = torch.cat([data.batch, torch.full((P,), B, dtype=data.batch.dtype, device=dev)])
        ptr = data.ptr.to(dev)
        out["ptr"] = torch.cat([ptr, torch.full((1,), N_b, dtype=ptr.dtype, device=dev)])
        out["cu_seqlens"] = torch.tensor([0, N, N_b], dtype=torch.int32).to(dev, non_blocking=True)
        out["valid_rows"] = torch.tensor([N], dtype=torch.int64).to(dev, non_blocking=True)
        if forces:      # energy per graph / force per node, pad entries zero (replaced inside the graph)
            out["energy"] = torch.cat([data.energy.reshape(-1).float(), torch.zeros(1, device=dev)])
            out["forces"] = torch.cat([data.forces.float(), torch.zeros((P, 3), device=dev)])
        return out

    # -------------------------------------------------------------------------- the step body ---
    def _masked_loss(self, pred, static):
        """core.loss over the padded batch: pad rows carry their own (detached) prediction as the target
        and the mean is rescaled to the real count."""
        core = self.core
        fn = core.loss_function
        total, tasks = 0, []
        for ihead, (p, kind) in enumerate(zip(pred, core.head_type)):
            n_real = static["B"] if kind == "graph" else static["valid_rows"]
            rows = p.size(0)
            keep = torch.arange(rows, device=p.device).unsqueeze(1) < n_real.view(1, 1)
            target = static["targets"][ihead]
            y = torch.where(keep, target, p.detach())
            scale = rows / n_real.to(p.dtype)
            term = fn(p, y) * scale.squeeze()
            total = total + term * core.loss_weights[ihead]
            tasks.append(term)
        return total, tasks

    def _masked_force_loss(self, pred, static):
        """Base.energy_force_loss (Base.py:582-636 of the reference) over the padded batch: the pad graph's
        energy and the pad nodes' forces are replaced by their own detached predictions, means rescaled."""
        from ..graph import pool_topo
        from ..ops import segment_sum

        core, data = self.core, static["data"]
        fn = core.loss_function
        n_graphs, n_nodes = static["B"].to(torch.float32), static["valid_rows"].to(torch.float32)
        e_pred = segment_sum(pred[0], pool_topo(data, pred[0].size(0)), 0).squeeze(-1).float()       # [B + 1]
        rows_g, rows_n = e_pred.size(0), data.pos.size(0)
        keep_g = torch.arange(rows_g, device=e_pred.device) < static["B"]
        e_true = static["energy"]
        w_e = core.loss_weights[0]
        energy_loss = fn(e_pred, torch.where(keep_g, e_true, e_pred.detach())) * (rows_g / n_graphs).squeeze()
        f_pred = -torch.autograd.grad(e_pred, data.pos, grad_outputs=torch.ones_like(e_pred), retain_graph=True,
                                      create_graph=True)[0].float()
        keep_n = (torch.arange(rows_n, device=e_pred.device) < static["valid_rows"]).unsqueeze(1)
        f_true = static["forces"]
        force_loss = fn(f_pred, torch.where(keep_n, f_true, f_pred.detach())) * (rows_n / n_nodes).squeeze()
        # weight: mean |E_true| / (mean |F_true| + 1e-8) over the real entries (pad entries are zero)
        w_f = w_e * (e_true.abs().sum() / n_graphs.squeeze()) / (f_true.abs().sum() / (3.0 * n_nodes.squeeze()) + 1e-8)
        return energy_loss * w_e + force_loss * w_f, [energy_loss]

    def _body(self, static):
        from ..globalAtt import graphed_tail
        data = static["data"]
        node_ops.valid_rows = static["valid_rows"]
        try:
            with graphed_tail.no_capture():       # no graphs of their own inside the step graph
                pred = self.model(data)
        finally:
            node_ops.valid_rows = None
        if static["forces"] is not None:
            loss, tasks = self._masked_force_loss(pred, static)
        else:
            loss, tasks = self._masked_loss(pred, static)
        loss.backward()
        return loss.detach(), [t.detach() for t in tasks]

    # ------------------------------------------------------------------------- gradients ---
    def _discover(self, static):
        """Which parameters this model's backward reaches (an eager optimizer skips the others: no weight
        decay, no moment update) -- one backward with .grad = None.  The flat buffer covers exactly those."""
        params = [p for p in self.core.parameters() if p.requires_grad]
        for p in params:
            p.grad = None
        if static["data"].pos.requires_grad:
            static["data"].pos.grad = None
        self._clear_caches(static)
        self._body(static)
        used = [p for p in params if p.grad is not None]
        if self._views is None:
            total = sum(p.numel() for p in used)
            self._flat = torch.zeros(total, dtype=torch.float32, device=used[0].device)
            self._views, off = [], 0
            for p in used:
                self._views.append((p, self._flat[off:off + p.numel()].view_as(p)))
                off += p.numel()
        elif [id(p) for p in used] != [id(p) for p, _ in self._views]:
            raise RuntimeError("step graph: this bucket's backward reaches a different set of parameters than the "
                               "first one did; the captured optimizer step cannot serve both")

    def _bind_grads(self):
        """p.grad = the parameter's view of the flat buffer (parameters backward never reaches: None)."""
        for p, v in self._views:
            p.grad = v
        averager = getattr(self.model, "averager", None)
        if averager is not None:
            averager.suspended = True           # its per-chunk hooks stay quiet: the mean is taken here

    def _fwd_bwd(self, static):
        """Forward, loss, backward; the gradients land in the flat buffer by ONE concatenation (backward is
        handed .grad = None, so there is no accumulate kernel per parameter), .grad = views of the buffer."""
        for p, _ in self._views:
            p.grad = None
        if static["data"].pos.requires_grad:
            static["data"].pos.grad = None
        out = self._body(static)
        torch.cat([p.grad.reshape(-1) for p, _ in self._views], out=self._flat)
        self._bind_grads()
        return out

    def _reduce_and_step(self):
        if self._world > 1:
            import torch.distributed as dist
            op = dist.ReduceOp.AVG if dist.get_backend() == "nccl" else dist.ReduceOp.SUM
            dist.all_reduce(self._flat, op=op)
            if op == dist.ReduceOp.SUM:
                self._flat.mul_(1.0 / self._world)
        self._opt_graph.replay()

    # ----------------------------------------------------------------------------- targets ---
    def _targets(self, data, padded, key):
        """Per head, the target rows in the padded layout ([B + 1, d] for graph heads, [N_b, d] for node
        heads), real rows first; pad rows zero (they are replaced inside the graph)."""
        core = self.core
        N_b, _, B, forces = key
        if forces:
            return []
        head_index = get_head_indices(self.model, data)
        out = []
        for ihead, kind in enumerate(core.head_type):
            d = core.head_dims[ihead]
            y = data.y[head_index[ihead]].reshape(-1, d)
            rows = (B + 1) if kind == "graph" else N_b
            full = torch.zeros((rows, d), dtype=y.dtype, device=y.device)
            full[:y.size(0)] = y
            out.append(full)
        return out

    # ------------------------------------------------------------------------------- capture ---
    def _make_static(self, data, padded, key):
        N_b, E_b, B, forces = key
        fields = {k: v.clone() for k, v in padded.items() if k in _NODE + _EDGE + ("edge_index", "batch", "ptr")}
        fields["num_graphs"] = B + 1
        sd = Data(**fields)
        sd.gps_cu_seqlens = padded["cu_seqlens"].clone()
        sd.gps_cu_seqlens._hgk_sum_sq = N_b * N_b
        sd.gps_cu_seqlens._hgk_bounds = None
        static = {"data": sd, "valid_rows": padded["valid_rows"].clone(),
                  "B": torch.tensor([B], dtype=torch.int64, device=sd.pos.device),
                  "targets": [t.clone() for t in self._targets(data, padded, key)],
                  "energy": padded["energy"].clone() if forces else None,
                  "forces": padded["forces"].clone() if forces else None}
        if forces:
            sd.pos.requires_grad_(True)        # the static position buffer is the leaf the forces are taken at
        if self._own_graph():
            # the in-model RadiusInteractionGraph takes the graph from this cache instead of searching
            sd.pos._hgk_rgraph_static = sd.edge_index
        return static

    def _fill(self, static, data, padded, key):
        sd = static["data"]
        with torch.no_grad():
            for k in _NODE + _EDGE + ("edge_index", "batch", "ptr"):
                if k in padded and getattr(sd, k, None) is not None:
                    getattr(sd, k).copy_(padded[k], non_blocking=True)
            if static["forces"] is not None:
                static["energy"].copy_(padded["energy"], non_blocking=True)
                static["forces"].copy_(padded["forces"], non_blocking=True)
        sd.gps_cu_seqlens.copy_(padded["cu_seqlens"], non_blocking=True)
        static["valid_rows"].copy_(padded["valid_rows"], non_blocking=True)
        for dst, src in zip(static["targets"], self._targets(data, padded, key)):
            dst.copy_(src, non_blocking=True)

    # The warm-up runs before a capture are real optimisation steps; model and optimizer state are put
    # back IN PLACE afterwards (graphs captured earlier hold the addresses of these tensors).
    def _snapshot(self):
        model = [(t, t.detach().clone()) for t in list(self.core.parameters()) + list(self.core.buffers())]
        opt = {}
        for p, st in self.optimizer.state.items():
            opt[p] = {k: v.detach().clone() for k, v in st.items() if torch.is_tensor(v)}
        return model, opt

    def _restore(self, snap):
        model, opt = snap
        with torch.no_grad():
            for t, saved in model:
                t.copy_(saved)
            for p, st in self.optimizer.state.items():
                for k, v in st.items():
                    if torch.is_tensor(v):
                        if p in opt and k in opt[p]:
                            v.copy_(opt[p][k])
                        else:
                            v.zero_()          # state created by the warm-up itself: back to "never stepped"

    @staticmethod
    def _clear_caches(static):
        """Topologies / CSRs cached on the static tensors must be rebuilt INSIDE the captured region."""
        sd = static["data"]
        for t in (sd.edge_index, sd.batch, sd.pos):
            for attr in ("_hgk_topo", "_hgk_ptr", "_hgk_rgraph"):
                if hasattr(t, attr):
                    try:
                        delattr(t, attr)
                    except AttributeError:
                        pass
        for attr in ("_pool_topo",):
            if hasattr(sd, attr):
                delattr(sd, attr)

    def _capture(self, data, padded, key):
        entry = _Entry()
        entry.static = self._make_static(data, padded, key)
        entry.uses = 0
        lr_dev = self._lr_tensors(data.pos.device)
        self._sync_lr()
        host_lr = [g["lr"] for g in self.optimizer.param_groups]
        for g, t in zip(self.optimizer.param_groups, lr_dev):
            g["lr"] = t
        try:
            return self._capture_with_tensor_lr(entry, data, padded, key)
        finally:
            for g, v in zip(self.optimizer.param_groups, host_lr):
                g["lr"] = v

    def _capture_with_tensor_lr(self, entry, data, padded, key):
        snap = self._snapshot()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            self._discover(entry.static)
            self._bind_grads()
            for _ in range(self.warmup):                 # eager runs: allocator / lazy-init warm-up, no collective
                self._clear_caches(entry.static)
                self._fwd_bwd(entry.static)
                self.optimizer.step()
        torch.cuda.current_stream().wait_stream(side)
        if self._opt_graph is None:                      # the optimizer step: one graph for all buckets
            opt_graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(opt_graph):
                self.optimizer.step()
            self._opt_graph = opt_graph
        self._restore(snap)
        self._clear_caches(entry.static)
        graph = torch.cuda.CUDAGraph()
        before = _hgk.stats.launches()
        with torch.cuda.graph(graph):
            entry.loss, entry.tasks = self._fwd_bwd(entry.static)
        entry.launches = _hgk.stats.launches() - before      # libhgk kernels inside one replay
        entry.graph = graph
        self.captures += 1
        return entry

    def _lr_tensors(self, device):
        if self._lr_dev is None:
            self._lr_host = [float(g["lr"]) for g in self.optimizer.param_groups]
            self._lr_dev = [torch.tensor(v, dtype=torch.float32, device=device) for v in self._lr_host]
        return self._lr_dev

    def _sync_lr(self):
        if self._lr_dev is None:
            return
        for k, g in enumerate(self.optimizer.param_groups):
            v = float(g["lr"])
            if v != self._lr_host[k]:
                self._lr_dev[k].fill_(v)
                self._lr_host[k] = v

    # ---------------------------------------------------------------------------------- step ---
    def eager_step(self, data, compute_grad_energy=False):
        """The plain `train_step` on the same model / optimizer (ineligible batches, instrumented runs)."""
        self.eager_steps += 1
        averager = getattr(self.model, "averager", None)
        if averager is not None:
            averager.suspended = False
        return train_step(self.model, self.optimizer, data, compute_grad_energy)

    def step(self, data, compute_grad_energy=False):
        """One optimisation step; returns (loss, per-task losses) like `train_step`."""
        if not self.eligible(data, compute_grad_energy):
            return self.eager_step(data, compute_grad_energy)
        key = self._key_of(data, compute_grad_energy)
        padded = self._pad(data, key)
        entry = self.entries.get(key)
        if entry is None:
            entry = self.entries[key] = self._capture(data, padded, key)
        self._fill(entry.static, data, padded, key)
        self._sync_lr()
        if self._views[0][0].grad is not self._views[0][1]:
            self._bind_grads()                   # somebody (an eager step's zero_grad) dropped the views
        entry.graph.replay()
        self._reduce_and_step()
        entry.uses += 1
        self.replays += 1
        _hgk.stats.replayed += entry.launches
        return entry.loss, entry.tasks
