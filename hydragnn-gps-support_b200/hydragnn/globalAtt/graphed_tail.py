"""CUDA-graph replay of the node-level parts of a GPS layer.

After the kernel work a training step is bounded by the host as much as by the GPU: the GPS block's
residual / BatchNorm / dropout / Linear chain is ~65 tiny launches per layer and step through eager
dispatch.  Everything in that chain is a function of [N, C] node matrices only -- its shapes depend
on the number of nodes of the mini-batch, not on the edges -- so for a given N the forward and the
backward of `GPSConv._tail_pre` / `_tail_post` can each be captured once as a CUDA graph
(torch.cuda.make_graphed_callables) and replayed as a single launch.  The attention kernel between
the two pieces stays an ordinary launch.  Datasets of fixed-size structures (the FeSi crystals of
BASELINE config 5: 1024 atoms per graph) present the same N every step; for a new N the eager path
runs until that N has been seen `CAPTURE_AFTER` times.

Not used (eager path instead): evaluation / no-grad, CPU tensors, and inside `eager()` -- the force
loss differentiates the backward pass itself (create_graph=True), which a captured backward cannot
provide; Base.forward enters it when the positions require grad.

Opt out with HYDRAGNN_GRAPH_TAIL=0.
"""
import contextlib
import os
import warnings

import torch

# make_graphed_callables captures its backward graphs on a side stream; autograd then notes, once per
# process, that the parameters' AccumulateGrad nodes live on another stream.  Harmless here (the
# replay is ordered after the producing stream), and noise in every training log.
warnings.filterwarnings("ignore", message=".*AccumulateGrad node's stream does not match.*")

ENABLED = os.environ.get("HYDRAGNN_GRAPH_TAIL", "1") != "0"
CAPTURE_AFTER = 3          # occurrences of one (layer, piece, N) signature before it is captured
MAX_GRAPHS_PER_LAYER = 8
_eager_depth = 0


@contextlib.contextmanager
def eager():
    """Run GPS layers eagerly inside this block AND on twice-differentiable torch ops (force loss: the
    backward pass is itself differentiated; GPSConv._native reads this flag)."""
    global _eager_depth
    _eager_depth += 1
    try:
        yield
    finally:
        _eager_depth -= 1


_no_capture_depth = 0


@contextlib.contextmanager
def no_capture():
    """No CUDA graphs of the tail's own inside this block (the caller captures the whole step itself,
    train/step_graph.py); the kernel-library tail stays selected."""
    global _no_capture_depth
    _no_capture_depth += 1
    try:
        yield
    finally:
        _no_capture_depth -= 1


class _Piece(torch.nn.Module):
    """`layer._tail_pre` / `_tail_post` as a module of tensors only.  It shares the layer's parameter
    objects and is deliberately NOT registered as a child of the layer (no extra state_dict keys)."""

    def __init__(self, layer, which, has_local, native=False):
        super().__init__()
        # register exactly the parameters the piece uses: make_graphed_callables differentiates the
        # captured backward with respect to every parameter of the module.  On the kernel-library path
        # (GPSConv._native) norm1 is applied in the second piece, together with norm2.
        if which == "pre":
            if not native:
                self.norm1 = layer.norm1
            self.in_w, self.in_b = layer.attn.in_proj_weight, layer.attn.in_proj_bias
        else:
            self.out_proj, self.mlp = layer.attn.out_proj, layer.mlp
            self.norm2, self.norm3 = layer.norm2, layer.norm3
            if native and has_local:
                self.norm1 = layer.norm1
        self._layer = [layer]
        self.which, self.has_local = which, has_local

    def forward(self, *args):
        layer = self._layer[0]
        if self.which == "pre":
            x, h_local = args
            return layer._tail_pre(x, h_local)
        if self.has_local:
            x, h1, o = args
            return layer._tail_post(x, h1, o)
        x, o = args
        return layer._tail_post(x, None, o)


def _batchnorm_state(layer):
    out = []
    for norm in (layer.norm1, layer.norm2, layer.norm3):
        if norm is None:
            continue
        for _, buf in norm.named_buffers():
            out.append((buf, buf.detach().clone()))
    return out


def _capture(layer, which, has_local, args, native=False):
    """Graph one piece for this input signature.  The warm-up passes of make_graphed_callables run
    real forwards: BatchNorm running statistics are restored afterwards (the random stream does
    advance -- dropout masks are not reproducible across devices anyway)."""
    saved = _batchnorm_state(layer)
    piece = _Piece(layer, which, has_local, native)
    piece.train()
    sample = tuple(a.detach().clone().requires_grad_(True) for a in args)
    graphed = torch.cuda.make_graphed_callables(piece, sample)
    for buf, value in saved:
        buf.copy_(value)
    return graphed


def _eager(layer, which, x, *rest):
    if which == "pre":
        return layer._tail_pre(x, rest[0])
    return layer._tail_post(x, rest[0], rest[1])


def run(layer, which, x, *rest):
    """`layer._tail_pre(x, h_local)` or `layer._tail_post(x, h1, o)`, replayed from a CUDA graph when
    this signature has been captured."""
    has_local = rest[0] is not None
    usable = (ENABLED and _eager_depth == 0 and _no_capture_depth == 0 and layer.training and torch.is_grad_enabled() and x.is_cuda
              and x.requires_grad and (which == "post" or has_local)
              and not torch.cuda.is_current_stream_capturing())
    if not usable:
        return _eager(layer, which, x, *rest)
    cache = layer.__dict__.setdefault("_graphed_tails", {})
    key = (which, x.size(0), x.size(1), has_local)
    entry = cache.get(key)
    if entry is None:
        entry = cache[key] = {"seen": 0, "graphed": None}
    args = (x,) + tuple(r for r in rest if r is not None)
    if entry["graphed"] is None:
        entry["seen"] += 1
        ready = sum(1 for e in cache.values() if e["graphed"] is not None)
        if entry["seen"] < CAPTURE_AFTER or ready >= MAX_GRAPHS_PER_LAYER or not all(a.requires_grad for a in args):
            return _eager(layer, which, x, *rest)
        native = bool(layer._native(x))
        if not all(p.requires_grad for p in _Piece(layer, which, has_local, native).parameters()):
            entry["seen"] = -(1 << 30)                 # frozen parameters (freeze_conv): never capture this one
            return _eager(layer, which, x, *rest)
        entry["graphed"] = _capture(layer, which, has_local, args, native)
    return entry["graphed"](*args)
