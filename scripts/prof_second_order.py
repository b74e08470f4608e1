"""Per-kernel CUDA-event times of the fused second-order tracking iteration at config C2 (un-graphed): python scripts/prof_second_order.py"""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from monogs_b200 import _lib as L
from monogs_b200.tracking import PoseTracker
dev = torch.device("cuda", 0)
tr, scd, cam, gt = bench.build_tracker("C2", dev, use_graph=False)
for _ in range(10):
    tr.step()
tr.begin_second_order(1e-3)
g = torch.Generator(device=dev).manual_seed(0)
idx, w = PoseTracker.gen_sketch_args(cam.H, cam.W, 16, 64, device=dev, generator=g)
for _ in range(3):
    tr.set_sketch(idx, w, 64); tr.second_order_step()
torch.cuda.synchronize()
lib = L.lib()
lib.mgs_profile_enable(1)
n = 20
for _ in range(n):
    tr.set_sketch(idx, w, 64); tr.second_order_step()
buf = C.create_string_buffer(8192)
lib.mgs_profile_report(buf, 8192)
lib.mgs_profile_enable(0)
tot = 0.0
for line in buf.value.decode().strip().splitlines():
    name, k, ms = line.split()
    print(f"{name:28s} {int(k):5d} launches  {1e3 * float(ms) / int(k):8.2f} us avg")
    tot += 1e3 * float(ms) / n
print("sum per iteration (us):", round(tot, 1))
