"""Per-kernel CUDA-event times of the fused tracking iteration (config C2 by default), un-graphed, library-side events
(mgs_profile_*): `python scripts/prof_events.py [N] [CFG]`.  Used to compare kernel variants (e.g. MGS_TRK_SPLIT=1|2)."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from monogs_b200 import _lib as L
n = int(sys.argv[1]) if len(sys.argv) > 1 else 50
cfg = sys.argv[2] if len(sys.argv) > 2 else "C2"
dev = torch.device("cuda", 0)
tr, scd, cam, gt = bench.build_tracker(cfg, dev, use_graph=False)
flush = torch.empty(256 << 20, device=dev, dtype=torch.uint8)
for _ in range(10):
    tr.step()
torch.cuda.synchronize()
lib = L.lib()
lib.mgs_profile_enable(1)
for _ in range(n):
    flush.zero_()
    tr.step()
buf = C.create_string_buffer(8192)
lib.mgs_profile_report(buf, 8192)
lib.mgs_profile_enable(0)
tot = 0.0
for line in buf.value.decode().strip().splitlines():
    name, k, ms = line.split()
    us = 1e3 * float(ms) / n
    tot += us
    print(f"{name:28s} {int(k) // n:3d} launch/iter {us:9.2f} us / iteration")
print(f"sum of kernels per iteration: {tot:.1f} us   (split={os.environ.get('MGS_TRK_SPLIT', 'default')}, cfg={cfg})")
# graph-replayed iteration time, L2 flushed between iterations
tr2, _, _, _ = bench.build_tracker(cfg, dev, use_graph=True)
ms = bench.timed_steps(tr2.step, 200, 20, flush, lambda: None)
print(f"graph replay: {ms / 200 * 1e3:.1f} us / iteration = {200 / ms * 1e3:.0f} it/s, loss {float(tr2.loss):.6f}")
