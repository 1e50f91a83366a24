"""tools/e2e_breakdown.py -- where the end-to-end step (bench.py `e2e`) spends its time beyond the kernel: host-side set-up of
sample_batch, the launch, the device-to-host reads."""
import os.path as osp
import sys
import time

import torch

sys.path.insert(0, osp.dirname(osp.dirname(osp.abspath(__file__))))
from bench import build_sampler  # noqa: E402

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
N = 1 << 20
env, policy, sampler = build_sampler("qq", torch, dev, 64, 1, 600)
for i in range(3):
    res = sampler.sample_batch(N, auto_reset=True)
    torch.cuda.synchronize()
rows = []
for i in range(5):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    res = sampler.sample_batch(N, auto_reset=True)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    a, b, c = res["ret"].cpu(), res["length"].cpu(), sampler.last_stats.cpu()
    t3 = time.perf_counter()
    rows.append((t1 - t0, t2 - t1, t3 - t2))
for r in rows:
    print("host set-up + launch %.2f ms | wait for the kernel %.2f ms | D2H reads %.2f ms" % tuple(1e3 * x for x in r))
import cProfile, pstats
pr = cProfile.Profile()
pr.enable()
res = sampler.sample_batch(N, auto_reset=True)
pr.disable()
torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
