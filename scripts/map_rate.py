"""Mapping iterations/s of config C4 on one GPU (bench.mapping_leg without the rest of the bench): python scripts/map_rate.py [window]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
window = int(sys.argv[1]) if len(sys.argv) > 1 else 8     # 1: the dependent chain of a rank that owns one view of the window
r = bench.mapping_leg(dev, 0, 1, iters=60, warmup=8, window=window)
print(f"mapping C4 window {window}: {r['iters_per_s']:.1f} it/s  (MGS_SCAN_FOLD={os.environ.get('MGS_SCAN_FOLD', '0')}, MGS_PDL={os.environ.get('MGS_PDL', '1')})")
