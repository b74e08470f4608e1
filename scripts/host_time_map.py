"""Host-side enqueue time vs device time of one MapEngine iteration (config C4): is the iteration launch-bound?"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from monogs_b200 import mapping as MP
dev = torch.device("cuda", 0)
orig = MP.MapEngine.step
host = []
def timed(self, views):
    t0 = time.perf_counter()
    r = orig(self, views)
    host.append(time.perf_counter() - t0)
    return r
MP.MapEngine.step = timed
window = int(sys.argv[1]) if len(sys.argv) > 1 else 8   # 1 = what a rank of an 8-GPU group enqueues per iteration
r = bench.mapping_leg(dev, 0, 1, iters=40, warmup=5, window=window)
h = sorted(host[5:])
print("device it/s", r["iters_per_s"], "-> ms/iter", 1e3 / r["iters_per_s"], " host enqueue ms/iter median", 1e3 * h[len(h) // 2])
