"""Host-side cost of one small-batch training step through the public module API (cProfile, B200 box).
usage: python scripts/profile_host.py [c2|c1r] [log2_points]"""
import cProfile
import io
import os
import pstats
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import WORKLOADS, Job  # noqa: E402

wl = WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "c2"]
n = 2 ** int(sys.argv[2] if len(sys.argv) > 2 else 12)
dev = torch.device("cuda:0")
job = Job(wl, dev, seed=1234, n_points=n)
params = list(job.enc.parameters())


def step():
    for p in params:
        p.grad = None
    res = job.enc.encode_features(job.coords)
    res.backward(job.gout)


for _ in range(20):
    step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(200):
    step()
t_host = time.perf_counter() - t0
torch.cuda.synchronize()
t_all = time.perf_counter() - t0
print(f"{sys.argv[1:]}: host issue {t_host / 200 * 1e3:.3f} ms/step, with device drain {t_all / 200 * 1e3:.3f} ms/step")
pr = cProfile.Profile()
pr.enable()
for _ in range(200):
    step()
pr.disable()
torch.cuda.synchronize()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(22)
print(s.getvalue()[:6000])
