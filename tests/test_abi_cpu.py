"""CPU: libhgk.so builds, loads, and exports every symbol include/hgk.h declares (no compute)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "hgk.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(hgk_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(libhgk):
    from hydragnn import _hgk

    names = declared_symbols()
    assert len(names) >= 15
    handle = ctypes.CDLL(_hgk.LIB_PATH)
    for n in names:
        assert hasattr(handle, n), f"{n} declared in hgk.h but not exported"
    assert set(names) == set(_hgk.EXPORTED_SYMBOLS), set(names) ^ set(_hgk.EXPORTED_SYMBOLS)


def test_argument_errors_are_reported_without_a_gpu(libhgk):
    assert libhgk.hgk_version() >= 100
    assert libhgk.hgk_csr_build_workspace_bytes(10, 4) > 0
    rc = libhgk.hgk_csr_build(None, -1, 4, None, None, None, 0, None)
    assert rc == -1 and b"negative" in libhgk.hgk_last_error()


def test_product_refuses_cpu_tensors(libhgk):
    import pytest
    import torch
    from hydragnn import _hgk
    from hydragnn.ops import Topo, segment_sum

    topo = Topo(torch.zeros(3, dtype=torch.int32), None, 2)
    with pytest.raises((_hgk.HgkError, RuntimeError)):
        segment_sum(torch.ones(3, 4), topo, 0)
