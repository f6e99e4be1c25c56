"""Developer run: the C3 prefix-scoring walk of bench.py (target of ncu launch lists)."""
import sys, time, torch
sys.path.insert(0, ".")
import bench
torch.cuda.set_device(0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20
t0 = time.perf_counter()
out = bench.bench_prefix(torch.device("cuda", 0), steps=n) if "--ref" in sys.argv else None
if out is None:
    import oracle.prefix_th_port as _p
    _p.PrefixScoreEager = None  # skip the eager reference: only our path under the profiler
    try:
        out = bench.bench_prefix(torch.device("cuda", 0), steps=n)
    except Exception as e:
        print("reference skipped:", type(e).__name__)
print(out["value"] if out else None, "steps/s; wall", time.perf_counter() - t0)
