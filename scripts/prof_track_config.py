"""Per-kernel CUDA-event times of the fused tracking iteration at a BASELINE config: python scripts/prof_track_config.py C3"""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from monogs_b200 import _lib as L
cfg = sys.argv[1] if len(sys.argv) > 1 else "C3"
tr, scd, cam, gt = bench.build_tracker(cfg, torch.device("cuda", 0), use_graph=False)
for _ in range(5): tr.step()
torch.cuda.synchronize()
L.lib().mgs_profile_enable(1)
n = 20
for _ in range(n): tr.step()
torch.cuda.synchronize()
buf = C.create_string_buffer(8192)
L.lib().mgs_profile_report(buf, 8192)
tot = 0
for line in buf.value.decode().strip().splitlines():
    name, k, ms = line.split()
    print(f"{name:28s} {int(k):5d} launches  {1e3 * float(ms) / int(k):8.2f} us avg")
    tot += 1e3 * float(ms) / n
print("sum per iteration (us):", tot, " instances:", tr.check_capacity())
