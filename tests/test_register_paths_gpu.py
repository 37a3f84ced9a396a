"""`-m gpu`: the register-staged variants of the TMA-fed kernels (what runs when the driver will not encode a
tensor map, and the A/B switches HGK_ATTN_BWD_TMA / HGK_WGRAD_TMA / HGK_LINEAR_TMA / HGK_FILTER_BWD_TMA / HGK_FILTER_FWD_TMA = 0)
against the same oracle comparisons as the default paths.  The switches are read once per process, so the kernel
tests are re-run in a child process with the switches set."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))


def test_kernel_parity_with_tma_switched_off(libhgk):
    env = dict(os.environ, HGK_ATTN_BWD_TMA="0", HGK_WGRAD_TMA="0", HGK_FILTER_BWD_TMA="0", HGK_FILTER_FWD_TMA="0",
               HGK_LINEAR_TMA="0")
    cmd = [sys.executable, "-m", "pytest", os.path.join(HERE, "test_kernels_gpu.py"), os.path.join(HERE, "test_gps_tail.py"),
           "-q", "-x", "-m", "gpu", "-p", "no:cacheprovider",
           "-k", "attn_varlen or attn_packed or filter or cfconv or wgrad or linear_tc"]
    out = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=900, cwd=os.path.dirname(HERE))
    tail = (out.stdout + out.stderr)[-3000:]
    assert out.returncode == 0, tail
    assert " passed" in out.stdout and "failed" not in out.stdout, tail
