"""`-m gpu`: hgk_csr_build_static, the entry point for callers that are not allowed to allocate
device memory (their GPU memory belongs to another allocator, or they sit in an enqueue path where
cudaMalloc would stall the device).  It must group edges exactly like hgk_csr_build, which gets its
workspace from the caller, and it must do so call after call although every call reuses the same
built-in scratch array.  Collected last on purpose (file name), so a `-x` run has already executed
every other test when it gets here."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
MAX_EDGES, MAX_ROWS = 1048576, 262144      # HGK_CSR_STATIC_MAX_* in include/hgk.h


def _static_csr(lib, key, num_rows):
    E = key.numel()
    rowptr = torch.empty(num_rows + 1, dtype=torch.int32, device=key.device)
    perm = torch.empty(max(E, 1), dtype=torch.int32, device=key.device)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.hgk_csr_build_static(ctypes.c_void_p(key.data_ptr()), E, num_rows,
                                  ctypes.c_void_p(rowptr.data_ptr()), ctypes.c_void_p(perm.data_ptr()), stream)
    assert rc == 0, lib.hgk_last_error()
    return rowptr, perm[:E]


def test_static_scratch_csr_matches_workspace_csr_twice_in_a_row(libhgk):
    from hydragnn.ops.topo import build_csr

    gen = torch.Generator().manual_seed(2024)
    # second call is smaller than the first, so stale scratch contents from call 1 lie beyond
    # (and inside) what call 2 uses; third call is an edge-free graph
    for num_rows, E in [(8192, 474078), (700, 9000), (5, 0)]:
        key = torch.randint(0, num_rows, (E,), generator=gen, dtype=torch.int32).to(DEV)
        want_rowptr, want_perm = build_csr(key, num_rows)
        got_rowptr, got_perm = _static_csr(libhgk, key, num_rows)
        torch.cuda.synchronize()
        assert torch.equal(got_rowptr, want_rowptr)         # integer work: bit-exact
        assert torch.equal(got_perm, want_perm)


def test_static_scratch_rejects_sizes_beyond_the_built_in_array(libhgk):
    dummy = torch.zeros(4, dtype=torch.int32, device=DEV)
    ptr = ctypes.c_void_p(dummy.data_ptr())
    for E, N in [(MAX_EDGES + 1, 10), (10, MAX_ROWS + 1)]:
        rc = libhgk.hgk_csr_build_static(ptr, E, N, ptr, ptr, None)
        assert rc == -2                                       # HGK_ERR_UNSUPPORTED
        assert b"exceeds the built-in scratch" in libhgk.hgk_last_error()
