"""Developer check: where the host time of one plugin step (CTC.loss_fn + backward) goes (run on the GPU box)."""
import cProfile
import os
import pstats
import sys
import time

import torch

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from espnet_b200.ctc import CTC  # noqa: E402

cfg = bench.CONFIGS["c2"]
batch = bench.make_batch(cfg, 1236)
pool = bench.label_pool(cfg, 1236, batch["ilen"])
x = batch["x"].cuda().requires_grad_()
ctc = CTC(cfg["V"], 8, ctc_type="b200")
ilen = batch["ilen"]


def step(i):
    tg, olen = pool[i % len(pool)]
    x.grad = None
    loss = ctc.loss_fn(x.transpose(0, 1), tg, ilen, olen)
    loss.backward()
    return loss


for i in range(20):
    step(i)
torch.cuda.synchronize()
for n in (1, 2):
    t0 = time.perf_counter()
    for i in range(300):
        step(i)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"host issue {1e6 * (t1 - t0) / 300:.1f} us/step, with drain {1e6 * (t2 - t0) / 300:.1f} us/step")
# host cost alone: synchronise every step so that the GPU never back-pressures the launch queue
t = 0.0
for i in range(200):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    step(i)
    t += time.perf_counter() - t0
print(f"host only {1e6 * t / 200:.1f} us/step")
pr = cProfile.Profile()
pr.enable()
for i in range(300):
    torch.cuda.synchronize()
    step(i)
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
