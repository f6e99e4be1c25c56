import sys, json, torch
sys.path.insert(0, "/root/repo")
import bench
dev = torch.device("cuda", 0)
for kw in ({}, {"margin": 40}, {"pre_beam": 15}):
    print(json.dumps(bench.bench_decode_loop(dev, **kw)))
