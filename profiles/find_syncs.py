"""List the host<->device synchronisation points of a training step (torch.cuda.set_sync_debug_mode):
every warning is an op that blocked the host until the stream drained.
    python profiles/find_syncs.py [workload]"""
import os
import sys
import warnings

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
    name = sys.argv[1] if len(sys.argv) > 1 else "schnet_gps_fesi1024"
    dev = torch.device("cuda:0")
    spec = syn.WORKLOADS[name]
    host = bench.build_batch_gpu(name, syn.sample_graphs(name, spec["graphs_per_gpu"]), dev)
    model = get_distributed_model(create_model(**spec["arch"]))
    opt = torch.optim.AdamW(model.module.parameters(), lr=1e-3)
    resident = bench.to_data(host, dev, Data)

    def step():
        d = Data(**{k: (v.detach().clone() if k in ("edge_index", "pos") else v) for k, v in resident.__dict__.items()})
        train_step(model, opt, d, compute_grad_energy=spec["forces"])

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    torch.cuda.set_sync_debug_mode("warn")
    with warnings.catch_warnings(record=True) as caught:
        warnings.simplefilter("always")
        import traceback
        orig = warnings.showwarning
        sites = {}

        def show(message, category, filename, lineno, file=None, line=None):
            stack = [f for f in traceback.extract_stack() if "hydragnn" in f.filename or "bench" in f.filename]
            key = " <- ".join(f"{os.path.basename(f.filename)}:{f.lineno}" for f in reversed(stack[-4:]))
            sites[key] = sites.get(key, 0) + 1
        warnings.showwarning = show
        warnings.simplefilter("always")
    # record=True swallows showwarning; run again with the hook installed
    warnings.showwarning = show
    warnings.simplefilter("always")
    step()
    torch.cuda.set_sync_debug_mode("default")
    warnings.showwarning = orig
    print(f"# {name}: synchronising calls in one training step")
    for k, v in sorted(sites.items(), key=lambda kv: -kv[1]):
        print(f"{v:4d}x  {k}")
    if not sites:
        print("none")


if __name__ == "__main__":
    main()
