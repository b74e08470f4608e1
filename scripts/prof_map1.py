"""Per-kernel CUDA-event times of a ONE-view mapping iteration (the dependent chain of a rank that owns one view of the
window) at config C4: python scripts/prof_map1.py"""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from monogs_b200 import _lib as L
torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
bench.mapping_leg(dev, 0, 1, iters=2, warmup=2, window=1, use_graph=False)
L.lib().mgs_profile_enable(1)
n = 20
r = bench.mapping_leg(dev, 0, 1, iters=n, warmup=0, window=1, use_graph=False)
buf = C.create_string_buffer(8192)
L.lib().mgs_profile_report(buf, 8192)
L.lib().mgs_profile_enable(0)
tot = 0
for line in buf.value.decode().strip().splitlines():
    name, k, ms = line.split()
    print(f"{name:28s} {int(k):5d} launches  {1e3 * float(ms) / int(k):8.2f} us avg")
    tot += 1e3 * float(ms) / int(k)
print("sum (us):", round(tot, 1), " it/s un-graphed with events:", round(r["iters_per_s"], 1))
