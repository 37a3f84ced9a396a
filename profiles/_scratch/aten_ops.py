"""Which ATen ops (not libhgk launches) are left in one config-5 training step: count and device time per op name."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "hydragnn-gps-support_b200"))
import bench
from hydragnn.graph import Data
from hydragnn.models import create_model
from hydragnn.train import train_step
from hydragnn.utils import synthetic as syn
from hydragnn.utils.distributed import get_distributed_model
name = "schnet_gps_fesi1024"; dev = torch.device("cuda:0")
torch.backends.cuda.matmul.allow_tf32 = False
spec = syn.WORKLOADS[name]
host = bench.build_batch_gpu(name, syn.sample_graphs(name, spec["graphs_per_gpu"]), dev)
model = get_distributed_model(create_model(**spec["arch"]))
opt = torch.optim.AdamW(model.module.parameters(), lr=1e-3, fused=True)
res = bench.to_data(host, dev, Data)
def step():
    d = Data(**{k: (v.detach().clone() if k in ("edge_index", "pos") else v) for k, v in res.__dict__.items()})
    train_step(model, opt, d, compute_grad_energy=spec["forces"])
for _ in range(4): step()
torch.cuda.synchronize()
with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA, torch.profiler.ProfilerActivity.CPU], with_stack=False) as prof:
    step(); torch.cuda.synchronize()
rows = []
for e in prof.key_averages():
    t = getattr(e, "self_device_time_total", 0) or getattr(e, "self_cuda_time_total", 0)
    if e.key.startswith("aten::") and t > 0:
        rows.append((t, e.count, e.key))
rows.sort(reverse=True)
print(f"ATen ops with device time: {sum(r[0] for r in rows):.0f} us in {sum(r[1] for r in rows)} calls")
for r in rows[:30]: print(f"{r[0]:8.1f} us {r[1]:4d}x {r[2]}")
