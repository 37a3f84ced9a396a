"""ctypes binding of libhgk.so (the C ABI declared in include/hgk.h).

This is the only module that touches the shared library.  There is no CPU
fallback: if the library is missing or a tensor is not a CUDA tensor the call
raises.  Everything is launched on torch's current CUDA stream.
"""
import ctypes
import os

import torch

_PKG_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(_PKG_ROOT, "libhgk.so")

_c_i32 = ctypes.c_int32
_c_i64 = ctypes.c_int64
_c_ptr = ctypes.c_void_p
_c_f32 = ctypes.c_float
_c_f64 = ctypes.c_double

# name -> (restype, argtypes); mirrors include/hgk.h one to one
_SIGNATURES = {
    "hgk_version": (_c_i32, []),
    "hgk_last_error": (ctypes.c_char_p, []),
    "hgk_csr_build_workspace_bytes": (_c_i64, [_c_i64, _c_i64]),
    "hgk_csr_build": (_c_i32, [_c_ptr, _c_i64, _c_i64, _c_ptr, _c_ptr, _c_ptr, _c_i64, _c_ptr]),
    "hgk_csr_build_static": (_c_i32, [_c_ptr, _c_i64, _c_i64, _c_ptr, _c_ptr, _c_ptr]),
    "hgk_contract_node": (
        _c_i32,
        [_c_ptr, _c_i64, _c_i32, _c_ptr, _c_i64, _c_i32, _c_ptr, _c_i64, _c_ptr,
         _c_ptr, _c_ptr, _c_ptr, _c_i64, _c_i64, _c_i32, _c_i32, _c_ptr],
    ),
    "hgk_contract_edge_w": (
        _c_i32,
        [_c_ptr, _c_i64, _c_ptr, _c_i64, _c_i32, _c_ptr, _c_ptr, _c_i64, _c_i32,
         _c_ptr, _c_ptr, _c_i64, _c_i32, _c_i32, _c_ptr],
    ),
    "hgk_contract_edge_u": (
        _c_i32,
        [_c_ptr, _c_ptr, _c_i64, _c_i32, _c_ptr, _c_ptr, _c_i64, _c_i32, _c_ptr,
         _c_ptr, _c_i64, _c_i64, _c_i32, _c_i32, _c_ptr],
    ),
    "hgk_edge_gather_add": (
        _c_i32,
        [_c_ptr, _c_i64, _c_ptr, _c_i64, _c_ptr, _c_ptr, _c_i64, _c_ptr, _c_i64, _c_i32, _c_ptr],
    ),
    "hgk_schnet_filter_fwd": (
        _c_i32,
        [_c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_i64, _c_i32, _c_ptr, _c_i32, _c_f32, _c_f32,
         _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_i64, _c_i32, _c_ptr],
    ),
    "hgk_schnet_filter_bwd": (_c_i32, [_c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_f32, _c_ptr, _c_i64, _c_i32, _c_ptr]),
    "hgk_cfconv_filter_bwd": (_c_i32, [_c_ptr, _c_ptr, _c_i64, _c_ptr, _c_ptr, _c_i64, _c_ptr, _c_ptr, _c_ptr, _c_f32,
                                       _c_ptr, _c_i64, _c_i32, _c_ptr]),
    "hgk_cfconv_wgrad": (_c_i32, [_c_ptr, _c_ptr, _c_ptr, _c_i64, _c_ptr, _c_ptr, _c_i64, _c_ptr, _c_ptr, _c_f32,
                                  _c_ptr, _c_i64, _c_i64, _c_i32, _c_ptr]),
    "hgk_wgrad_tn_partials": (_c_i64, [_c_i64]),
    "hgk_wgrad_tn": (
        _c_i32,
        [_c_ptr, _c_ptr, _c_ptr, _c_i64, _c_ptr, _c_f32, _c_ptr, _c_i64, _c_i32, _c_ptr, _c_ptr, _c_i64, _c_i32,
         _c_ptr, _c_i32, _c_f32, _c_i64, _c_i32, _c_ptr],
    ),
    "hgk_colsum_slabs": (_c_i64, [_c_i64]),
    "hgk_colsum": (_c_i32, [_c_ptr, _c_ptr, _c_ptr, _c_i64, _c_i64, _c_i32, _c_ptr]),
    "hgk_wgrad_reduce": (_c_i32, [_c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_i64, _c_i32, _c_i32, _c_ptr]),
    "hgk_wgrad_rows_workspace_floats": (_c_i64, [_c_i64, _c_i32, _c_i32]),
    "hgk_wgrad_rows": (_c_i32, [_c_ptr, _c_i64, _c_ptr, _c_ptr, _c_ptr, _c_i64, _c_i32, _c_ptr, _c_i64, _c_i32, _c_i64, _c_ptr]),
    "hgk_linear_tc": (
        _c_i32,
        [_c_ptr, _c_i64, _c_ptr, _c_i64, _c_ptr, _c_i64, _c_i32, _c_ptr, _c_i32, _c_ptr, _c_ptr, _c_i64, _c_f32,
         _c_ptr, _c_i64, _c_i64, _c_i32, _c_i32, _c_ptr],
    ),
    "hgk_bn_slabs": (_c_i64, [_c_i64]),
    "hgk_bn_stats": (
        _c_i32,
        [_c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_f32, _c_f32, _c_ptr, _c_ptr, _c_i64, _c_i64, _c_i32, _c_ptr, _c_ptr],
    ),
    "hgk_bn_apply": (
        _c_i32,
        [_c_ptr, _c_i64, _c_ptr, _c_i64, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_i64, _c_ptr, _c_ptr, _c_ptr,
         _c_ptr, _c_i32, _c_i64, _c_i32, _c_ptr],
    ),
    "hgk_bn_bwd_workspace_floats": (_c_i64, [_c_i64, _c_i32]),
    "hgk_bn_bwd": (
        _c_i32,
        [_c_ptr, _c_ptr, _c_ptr, _c_f32, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_f32, _c_ptr, _c_ptr, _c_ptr,
         _c_ptr, _c_i64, _c_ptr, _c_i64, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_i64, _c_ptr, _c_ptr, _c_ptr,
         _c_i64, _c_i32, _c_ptr, _c_ptr],
    ),
    "hgk_mask_axpy": (_c_i32, [_c_ptr, _c_ptr, _c_i64, _c_ptr, _c_f32, _c_ptr, _c_i64, _c_i64, _c_i32, _c_ptr]),
    "hgk_edge_geom_fwd": (_c_i32, [_c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_f32, _c_i64, _c_ptr]),
    "hgk_edge_geom_bwd": (_c_i32, [_c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_f32, _c_i64, _c_ptr]),
    "hgk_edge_geom_bwd2": (_c_i32, [_c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_f32, _c_i64, _c_ptr]),
    "hgk_collate_rows": (_c_i32, [_c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_i64, _c_i32, _c_ptr, _c_i64, _c_ptr]),
    "hgk_collate_edges": (_c_i32, [_c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_i64, _c_i64, _c_ptr]),
    "hgk_pna_aggregate_fwd": (
        _c_i32,
        [_c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr,
         _c_i64, _c_i64, _c_i32, _c_ptr],
    ),
    "hgk_pna_aggregate_bwd": (
        _c_i32,
        [_c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr,
         _c_ptr, _c_ptr, _c_i64, _c_i64, _c_i32, _c_ptr],
    ),
    "hgk_radius_graph_workspace_bytes": (_c_i64, [_c_i64, _c_i64]),
    "hgk_radius_graph_count": (
        _c_i32,
        [_c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_i64, _c_i64, _c_f64, _c_i32, _c_i32, _c_i32,
         _c_ptr, _c_ptr, _c_ptr, _c_i64, _c_ptr],
    ),
    "hgk_radius_graph_fill": (
        _c_i32,
        [_c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_i64, _c_i64, _c_f64, _c_i32, _c_i32, _c_i32,
         _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_i64, _c_ptr, _c_i64, _c_ptr],
    ),
    "hgk_attn_varlen_fwd": (
        _c_i32,
        [_c_ptr, _c_i64, _c_ptr, _c_ptr, _c_i64, _c_ptr, _c_i64, _c_ptr, _c_i64,
         _c_ptr, _c_i64, _c_i64, _c_i32, _c_i32, _c_f32, _c_ptr],
    ),
    "hgk_attn_varlen_bwd": (
        _c_i32,
        [_c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_ptr, _c_i64, _c_ptr, _c_i64, _c_ptr,
         _c_ptr, _c_i64, _c_ptr, _c_i64, _c_ptr, _c_i64, _c_ptr, _c_i64, _c_i64,
         _c_i32, _c_i32, _c_f32, _c_ptr],
    ),
    "hgk_attn_varlen_bwd_ld": (
        _c_i32,
        [_c_ptr, _c_ptr, _c_ptr, _c_i64, _c_ptr, _c_ptr, _c_i64, _c_ptr, _c_i64, _c_ptr,
         _c_ptr, _c_i64, _c_ptr, _c_i64, _c_ptr, _c_i64, _c_ptr, _c_i64, _c_i64,
         _c_i32, _c_i32, _c_f32, _c_ptr],
    ),
}

EXPORTED_SYMBOLS = tuple(_SIGNATURES)

_lib = None


class HgkError(RuntimeError):
    pass


def lib():
    """Load libhgk.so once; raise loudly if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with "
                "`python hydragnn-gps-support_b200/csrc/build.py` "
                "(or __graft_entry__.build()). There is no CPU fallback."
            )
        # HGK_LIB: an instrumented build of the same sources (profiles/kv_phase_profile.py)
        handle = ctypes.CDLL(os.environ.get("HGK_LIB") or LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError if the symbol is absent
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(status, what):
    if status != 0:
        msg = lib().hgk_last_error()
        raise HgkError(f"{what} failed with status {status}: {msg.decode() if msg else ''}")


# CUDA kernels enqueued by one call of each entry point (for bench.py's gpu_launches count)
KERNELS_PER_CALL = {
    "hgk_csr_build": 4, "hgk_csr_build_static": 4, "hgk_contract_node": 1, "hgk_contract_edge_w": 1, "hgk_contract_edge_u": 1,
    "hgk_edge_gather_add": 1, "hgk_schnet_filter_fwd": 1, "hgk_schnet_filter_bwd": 1, "hgk_wgrad_tn": 1, "hgk_wgrad_reduce": 1, "hgk_colsum": 2, "hgk_pna_aggregate_fwd": 1, "hgk_pna_aggregate_bwd": 1,
    "hgk_radius_graph_count": 8, "hgk_radius_graph_fill": 1, "hgk_attn_varlen_fwd": 1,
    "hgk_attn_varlen_bwd": 3, "hgk_attn_varlen_bwd_ld": 3, "hgk_wgrad_rows": 2, "hgk_linear_tc": 1, "hgk_bn_stats": 2, "hgk_bn_apply": 1,
    "hgk_bn_bwd": 3, "hgk_mask_axpy": 1, "hgk_edge_geom_fwd": 1, "hgk_edge_geom_bwd": 1, "hgk_edge_geom_bwd2": 1,
    "hgk_collate_rows": 1, "hgk_collate_edges": 1, "hgk_cfconv_filter_bwd": 1, "hgk_cfconv_wgrad": 1,
}


class Stats:
    """Call counters and (optionally) CUDA-event timing of every C-ABI call.

    `cost` is the ALGORITHMIC work of the call -- (bytes, flops[, compulsory_bytes]) from the formulas
    in DESIGN.md -- so that achieved GB/s or TFLOP/s = sum(cost) / sum(event time) per entry point.
    `bytes` counts a gathered node row once per EDGE; `compulsory_bytes` counts the
    gathered table once (what must cross HBM when the table stays in the 126 MB L2)."""

    def __init__(self):
        self.reset()
        self.timing = False

    def reset(self):
        self.calls = {}
        self.cost = {}
        self.events = []
        self.replayed = 0           # libhgk kernels launched by CUDA-graph replays (train/step_graph.py)

    def launches(self):
        return sum(KERNELS_PER_CALL.get(k, 1) * v for k, v in self.calls.items()) + self.replayed

    def summary(self):
        """name -> dict(calls, ms, bytes, flops); call after torch.cuda.synchronize()."""
        out = {}
        for name, n in self.calls.items():
            b, f, d = self.cost.get(name, (0, 0, 0))
            out[name] = dict(calls=n, ms=0.0, bytes=b, flops=f, dram_bytes=d)
        for name, e0, e1 in self.events:
            out[name]["ms"] += e0.elapsed_time(e1)
        return out


stats = Stats()


def call(name, *args, cost=(0, 0)):
    """Invoke entry point `name` on torch's current stream; count it, optionally time it."""
    fn = getattr(lib(), name)
    stats.calls[name] = stats.calls.get(name, 0) + 1
    if cost != (0, 0):
        b, f, d = stats.cost.get(name, (0, 0, 0))
        stats.cost[name] = (b + cost[0], f + cost[1], d + (cost[2] if len(cost) > 2 else cost[0]))
    if stats.timing and not torch.cuda.is_current_stream_capturing():   # events recorded in a capture cannot be timed
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        status = fn(*args, stream_ptr())
        e1.record()
        stats.events.append((name, e0, e1))
    else:
        status = fn(*args, stream_ptr())
    check(status, name)


def stream_ptr():
    """Raw handle of torch's current stream on the current device (the cheap accessor: building a
    torch.cuda.Stream object per kernel call costs several microseconds of host time)."""
    return _c_ptr(torch._C._cuda_getCurrentRawStream(torch.cuda.current_device()))


def ptr(t):
    """Device pointer of a tensor (None -> NULL); refuses CPU tensors."""
    if t is None:
        return _c_ptr(0)
    if not t.is_cuda:
        raise HgkError("libhgk needs CUDA tensors; there is no CPU path")
    return _c_ptr(t.data_ptr())


def rows(t, channels=1):
    """Validate a [rows, F] or [rows, C, F] fp32 operand; return (tensor, ld).

    Rows may be strided (views into a wider matrix), columns must be unit
    stride and channels F apart."""
    if t.dtype != torch.float32:
        raise HgkError(f"expected float32, got {t.dtype}")
    if t.dim() == 2:
        if t.size(0) > 1 and t.size(1) > 0 and t.stride(1) != 1 and t.size(1) != 1:
            t = t.contiguous()
        elif t.size(1) == 1 and t.stride(1) != 1:
            t = t.contiguous()
        ld = t.stride(0) if t.size(0) > 1 else t.size(1)
        if ld < t.size(1):
            t = t.contiguous()
            ld = t.size(1)
        return t, ld
    if t.dim() == 3:
        F = t.size(2)
        ok = (t.stride(2) == 1 or F == 1) and (t.stride(1) == F or t.size(1) == 1)
        if not ok:
            t = t.contiguous()
        ld = t.stride(0) if t.size(0) > 1 else t.size(1) * F
        if ld < t.size(1) * F:
            t = t.contiguous()
            ld = t.size(1) * F
        return t, ld
    raise HgkError(f"expected a 2-D or 3-D tensor, got {t.dim()}-D")
