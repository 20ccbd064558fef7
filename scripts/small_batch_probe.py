import os, sys, time
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import torch
from bench import WORKLOADS, Job
dev = torch.device("cuda:0")
job = Job(WORKLOADS["c2"], dev, seed=1234, n_points=4096)
params = list(job.enc.parameters())
def step():
    for p in params: p.grad = None
    res = job.enc.encode_features(job.coords)
    res.backward(job.gout)
for _ in range(5): step()
torch.cuda.synchronize()
tabs = [p for p in params if p.dim() == 2 and p.shape[0] == 2 ** 19]
stor = {p.grad.untyped_storage().data_ptr() for p in tabs}
print("table grads:", len(tabs), "distinct storages:", len(stor))
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(3): step()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=14, max_name_column_width=60))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); e0.record()
for _ in range(100): step()
e1.record(); torch.cuda.synchronize()
print("ms/step (events)", e0.elapsed_time(e1) / 100)
