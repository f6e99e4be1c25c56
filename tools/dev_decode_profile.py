"""Developer run: which kernels / operators the C3 search loop (bench.bench_decode_loop) spends its time in."""
import sys, torch
sys.path.insert(0, ".")
import bench
from torch.profiler import profile, ProfilerActivity
dev = torch.device("cuda", 0)
kw = {"pre": {"pre_beam": 15}, "window": {"margin": 40}}.get(sys.argv[1] if len(sys.argv) > 1 else "", {})
bench.bench_decode_loop(dev, **kw)
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    out = bench.bench_decode_loop(dev, **kw)
print(out["us_per_position"], out["positions"])
print(prof.key_averages().table(sort_by="self_cuda_time_total", row_limit=18, max_name_column_width=60))
print(prof.key_averages().table(sort_by="self_cpu_time_total", row_limit=14, max_name_column_width=60))
