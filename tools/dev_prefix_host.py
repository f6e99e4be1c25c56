"""Developer run: where the host time of one prefix-scoring step goes (cProfile of the C3 pre-beam walk)."""
import cProfile, pstats, sys, torch
sys.path.insert(0, ".")
import bench
import oracle.stage_ref as sr
sr.activate = lambda: False
snum = int(sys.argv[1]) if len(sys.argv) > 1 else 15
dev = torch.device("cuda", 0)
bench.bench_prefix(dev, 6459.3, steps=20, snum=snum)
pr = cProfile.Profile()
pr.enable()
out = bench.bench_prefix(dev, 6459.3, steps=200, snum=snum)
pr.disable()
print(out["us_per_step"], "host", out["host_issue_us_per_step"])
pstats.Stats(pr).sort_stats("tottime").print_stats(18)
