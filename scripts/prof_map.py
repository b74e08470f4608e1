"""Short mapping run for ncu: a few iterations of the 8-keyframe mapping window (config C4) on one GPU."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2
torch.cuda.set_device(0)
r = bench.mapping_leg(torch.device("cuda", 0), 0, 1, iters=n, warmup=1)
print("ok", r["iters_per_s"])
