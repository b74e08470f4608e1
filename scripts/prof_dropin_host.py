"""Host-side profile (cProfile) of the drop-in tracking iteration that bench.py's e2e_dropin leg times.
    python scripts/prof_dropin_host.py [steps]"""
import cProfile, pstats, io, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from monogs_b200 import scene as S

n = int(sys.argv[1]) if len(sys.argv) > 1 else 300
dev = torch.device("cuda", 0)
sc, cam = S.make_config("C2")
gt_host = torch.rand(3, cam.H, cam.W).pin_memory()
print("rate: %.0f it/s" % bench.e2e_dropin_leg("C2", dev, n, 10, gt_host))
pr = cProfile.Profile()
pr.enable()
bench.e2e_dropin_leg("C2", dev, n, 10, gt_host)
pr.disable()
st = io.StringIO()
pstats.Stats(pr, stream=st).sort_stats("tottime").print_stats(45)
print(st.getvalue()[:9000])
