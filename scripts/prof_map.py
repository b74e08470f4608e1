"""Short mapping run: a few iterations of the 8-keyframe mapping window (config C4) on one GPU through MapEngine.
`python scripts/prof_map.py N` for ncu; `python scripts/prof_map.py N events` prints per-kernel CUDA-event times."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from monogs_b200 import _lib as L
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2
torch.cuda.set_device(0)
if len(sys.argv) > 2:
    bench.mapping_leg(torch.device("cuda", 0), 0, 1, iters=2, warmup=2)
    L.lib().mgs_profile_enable(1)
r = bench.mapping_leg(torch.device("cuda", 0), 0, 1, iters=n, warmup=1)
if len(sys.argv) > 2:
    buf = C.create_string_buffer(8192)
    L.lib().mgs_profile_report(buf, 8192)
    tot = 0
    for line in buf.value.decode().strip().splitlines():
        name, k, ms = line.split()
        print(f"{name:28s} {int(k):5d} launches  {1e3 * float(ms) / int(k):8.2f} us avg  {1e3 * float(ms) / (n + 1):9.1f} us / iteration")
        tot += 1e3 * float(ms) / (n + 1)
    print("sum per iteration (us):", tot)
print("ok", r["iters_per_s"])
