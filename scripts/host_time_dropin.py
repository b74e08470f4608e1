"""cProfile of the drop-in tracking step (GaussianRasterizer autograd surface + fused loss / Adam / pose ops, config C2):
where the host time of `e2e_dropin` goes.  python scripts/host_time_dropin.py [iters]"""
import cProfile, pstats, sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
n = int(sys.argv[1]) if len(sys.argv) > 1 else 200
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
tr, scd, cam, gt = bench.build_tracker("C2", dev, use_graph=False)
gt_host = gt.cpu().pin_memory()
del tr
print("warm it/s", bench.e2e_dropin_leg("C2", dev, 50, 5, gt_host))
pr = cProfile.Profile()
pr.enable()
v = bench.e2e_dropin_leg("C2", dev, n, 5, gt_host)
pr.disable()
print("profiled it/s", v)
pstats.Stats(pr).sort_stats("cumulative").print_stats(38)
