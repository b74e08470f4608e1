"""Short un-graphed tracking run for ncu: a few iterations of PoseTracker on config C2."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4
tr, scd, cam, gt = bench.build_tracker("C2", torch.device("cuda", 0), use_graph=False)
for _ in range(n):
    tr.step()
torch.cuda.synchronize()
print("ok", float(tr.loss))
