"""Developer run: the C3 prefix-scoring walk with scoring_ids = top-15 (target of ncu launch lists)."""
import sys, torch
sys.path.insert(0, ".")
import bench
import oracle.stage_ref as sr
sr.activate = lambda: False          # our path only under the profiler
if len(sys.argv) > 3:
    from espnet_b200 import _lib
    _lib.load().b200_ctc_debug_prefix_fp32_score({"fp32": 1, "copies": 2}[sys.argv[3]])
out = bench.bench_prefix(torch.device("cuda", 0), 6459.3, steps=int(sys.argv[1]) if len(sys.argv) > 1 else 6, snum=int(sys.argv[2]) if len(sys.argv) > 2 else 15)
print(out["value"], out["us_per_step"], "host", out["host_issue_us_per_step"], out.get("column_copy"))
