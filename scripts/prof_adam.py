"""Fused Adam at C5 scale (2M Gaussians, 59 floats each): CUDA-event GB/s.  For ncu: -k regex:adam_kernel."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from monogs_b200 import ops
P = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
dev = torch.device("cuda", 0)
shapes = {"xyz": (P, 3), "f_dc": (P, 1, 3), "f_rest": (P, 15, 3), "opacity": (P, 1), "scaling": (P, 3), "rotation": (P, 4)}
ps = {k: torch.randn(*s, device=dev) for k, s in shapes.items()}
for p in ps.values():
    p.grad = torch.randn_like(p) * 1e-3
opt = ops.FusedAdam([{"params": [p], "lr": 1e-3} for p in ps.values()], eps=1e-15)
for _ in range(3): opt.step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): opt.step()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
n = sum(p.numel() for p in ps.values())
print(f"adam: {n} elements, {ms * 1e3:.1f} us, {28 * n / ms / 1e6:.1f} GB/s")
