"""mgs_peer_reduce_adam on ONE GPU with world = 1 (the rank owns the whole arena; loads and stores go through the same
TMA pipeline, to local memory): the local-HBM side of the kernel for ncu.  A multi-rank group can not run under ncu
(kernel replay serialises the ranks that wait for each other).  `python scripts/prof_peer.py [C4|C5] [iters]`"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from monogs_b200 import _lib as L
from monogs_b200.peer import PeerBlock
from scripts.check_peer_exchange import layout

cfg = sys.argv[1] if len(sys.argv) > 1 else "C5"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 5
P, M = {"C4": (200_000, 1), "C5": (2_000_000, 16)}[cfg]
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
lay, lrs = layout(P, M, dev)
n = lay.numel
blk = PeerBlock.local_group(n, 1, dev)[0]
blk.params.normal_(); blk.grads.normal_()
m, v = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
x = blk.exchange(m, v, lay.segment_ends, lrs)
lib = L.lib()
s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
flush = torch.empty(256 << 20, device=dev, dtype=torch.uint8)
ev = []
for i in range(1, iters + 4):
    flush.zero_()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    L.check(lib.mgs_peer_reduce_adam(C.byref(x), i, i, 0.9, 0.999, 1e-15, s), "mgs_peer_reduce_adam")
    b.record()
    ev.append((a, b))
torch.cuda.synchronize()
us = sorted(1e3 * a.elapsed_time(b) for a, b in ev[3:])
med = us[len(us) // 2]
print(f"{cfg}: arena {4 * n / 1e6:.1f} MB, world 1: {med:.1f} us per exchange, {28 * n / med / 1e3:.0f} GB/s of Adam traffic (28 B/element)")
