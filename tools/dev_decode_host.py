"""Developer run: cProfile of the C3 search loop (where the host time of a position goes)."""
import cProfile, pstats, sys, torch
sys.path.insert(0, ".")
import bench
dev = torch.device("cuda", 0)
kw = {"pre": {"pre_beam": 15}, "window": {"margin": 40}}.get(sys.argv[1] if len(sys.argv) > 1 else "", {})
bench.bench_decode_loop(dev, **kw)
pr = cProfile.Profile()
pr.enable()
out = bench.bench_decode_loop(dev, **kw)
pr.disable()
print(out["us_per_position"], out["positions"])
pstats.Stats(pr).sort_stats("tottime").print_stats(28)
