"""Kernel-time table of a few training steps of a bench workload (torch.profiler, CUDA activities).
    python profiles/torch_profile.py [workload] [steps]"""
import os
import sys

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
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
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

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA,
                                            torch.profiler.ProfilerActivity.CPU]) as prof:
        for _ in range(steps):
            step()
        torch.cuda.synchronize()
    rows = []
    for e in prof.key_averages():
        t = getattr(e, "device_time_total", 0) or getattr(e, "cuda_time_total", 0)
        if e.device_type == torch.autograd.DeviceType.CUDA and t > 0:
            rows.append((t / steps, e.count / steps, e.key))
    rows.sort(reverse=True)
    total = sum(r[0] for r in rows)
    print(f"# {name}: {total / 1e3:.2f} ms of kernel time per step, {sum(r[1] for r in rows):.0f} launches per step")
    for t, n, k in rows[:45]:
        print(f"{t / total * 100:6.2f}% {t:9.1f} us {n:6.1f}x  {k[:110]}")
    # host side: ATen ops / autograd nodes by self CPU time
    cpu = [(e.self_cpu_time_total / steps, e.count / steps, e.key) for e in prof.key_averages()
           if e.device_type == torch.autograd.DeviceType.CPU]
    cpu.sort(reverse=True)
    print(f"# host: {sum(c[0] for c in cpu) / 1e3:.2f} ms of self CPU time per step in {sum(c[1] for c in cpu):.0f} ops")
    for t, n, k in cpu[:60]:
        print(f"{t:9.1f} us {n:6.1f}x  {k[:100]}")


if __name__ == "__main__":
    main()
