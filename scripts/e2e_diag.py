"""Run-to-run spread of the host-bound tracking loops (bench.py's e2e legs): python scripts/e2e_diag.py [config] [runs]"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
runs = int(sys.argv[2]) if len(sys.argv) > 2 else 8
dev = torch.device("cuda", 0)
tr, scd, cam, gt = bench.build_tracker(cfg, dev, use_graph=True)
gt_host = gt.cpu().pin_memory()
loss_host = torch.zeros(2).pin_memory(); tau_host = torch.zeros(2, 6).pin_memory()
depth_host = tr.bench_gt_depth.cpu().pin_memory() if tr.bench_gt_depth is not None else None
tr.bind_host(gt_host, loss_host, tau_host, depth_host)
K = 200
def lagged():
    slot = tr.step_host_lagged()
    if slot is not None: return float(loss_host[slot])
def sync():
    tr.step_host(); return float(loss_host[0])
def device_only():
    tr.step()
for name, fn in (("device", device_only), ("lagged", lagged), ("sync", sync), ("lagged", lagged), ("sync", sync)):
    rates = []
    for _ in range(runs):
        for _ in range(10): fn()
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(K): fn()
        torch.cuda.synchronize(); rates.append(K / (time.perf_counter() - t0))
    print(name, " ".join(f"{r:.0f}" for r in rates))
