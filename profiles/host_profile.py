"""Host-side (Python) cost of a training step: cProfile over a few steps of a bench workload,
sorted by own time.  The small workloads are bound by this, not by the GPU.
    python profiles/host_profile.py [workload] [steps]"""
import cProfile
import os
import pstats
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "hydragnn-gps-support_b200"))
import bench  # noqa: E402
from hydragnn.graph import Data  # noqa: E402
from hydragnn.models import create_model  # noqa: E402
from hydragnn.train import train_step  # noqa: E402
from hydragnn.utils import synthetic as syn  # noqa: E402
from hydragnn.utils.distributed import get_distributed_model  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "schnet_qm9"
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    dev = torch.device("cuda:0")
    torch.backends.cuda.matmul.allow_tf32 = False
    spec = syn.WORKLOADS[name]
    host = bench.build_batch_gpu(name, syn.sample_graphs(name, spec["graphs_per_gpu"]), dev)
    model = get_distributed_model(create_model(**spec["arch"]))
    opt = torch.optim.AdamW(model.module.parameters(), lr=1e-3)
    resident = bench.to_data(host, dev, Data)

    def step():
        d = Data(**{k: (v.detach().clone() if k in ("edge_index", "pos") else v) for k, v in resident.__dict__.items()})
        train_step(model, opt, d, compute_grad_energy=spec["forces"])

    for _ in range(5):
        step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    t_host = time.perf_counter() - t0
    torch.cuda.synchronize()
    t_all = time.perf_counter() - t0
    print(f"# {name}: host enqueue {t_host / steps * 1e3:.2f} ms/step, with GPU drain {t_all / steps * 1e3:.2f} ms/step")
    prof = cProfile.Profile()
    prof.enable()
    for _ in range(steps):
        step()
    prof.disable()
    torch.cuda.synchronize()
    st = pstats.Stats(prof)
    st.sort_stats("tottime").print_stats(45)
    st.sort_stats("cumtime").print_stats(60)
    st.print_callers("torch.tensor")


if __name__ == "__main__":
    main()
