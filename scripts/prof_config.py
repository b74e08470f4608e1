"""Per-kernel CUDA-event times of one MapEngine iteration (1 view) at a BASELINE config: python scripts/prof_config.py C5"""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from monogs_b200 import scene as S, rasterizer as RZ, _lib as L
from monogs_b200.mapping import MapEngine, MapView
cfg = sys.argv[1] if len(sys.argv) > 1 else "C5"
n_views = int(sys.argv[2]) if len(sys.argv) > 2 else 1
dev = torch.device("cuda", 0)
sc, cam = S.make_config(cfg)
to_dev = lambda nt: type(nt)(*[x.to(dev) if isinstance(x, torch.Tensor) else x for x in nt])
scd, camd = to_dev(sc), to_dev(cam)
bg = torch.zeros(3, device=dev)
rs = RZ.GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, bg, 1.0, camd.viewmatrix, camd.projmatrix,
                                      camd.projmatrix_raw, sc.sh_degree, camd.campos, False, False)
gt, _, gt_depth, _, _, _ = RZ.raster_forward(rs, scd.means3D, scd.opacities, scd.shs, None, scd.scales, scd.rotations, None)
oc = scd.opacities.clamp(1e-4, 1 - 1e-4)
raw = {"xyz": scd.means3D.clone(), "f_dc": scd.shs[:, :1].contiguous(), "opacity": torch.log(oc / (1 - oc)),
       "scaling": torch.log(scd.scales), "rotation": scd.rotations.clone()}
if scd.shs.size(1) > 1: raw["f_rest"] = scd.shs[:, 1:].contiguous()
lrs = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 2.5e-3 / 20, "opacity": 0.05, "scaling": 1e-3, "rotation": 1e-3}
eng = MapEngine(raw, lrs, sh_degree=sc.sh_degree, W=cam.W, H=cam.H, capacity_headroom=1.5, overlap=False)
views = [MapView(rs, (gt * 0.9 + 0.05).contiguous(), None) for _ in range(n_views)]
for _ in range(3): eng.step(views)
torch.cuda.synchronize()
L.lib().mgs_profile_enable(1)
n = 5
for _ in range(n): eng.step(views)
torch.cuda.synchronize()
buf = C.create_string_buffer(8192)
L.lib().mgs_profile_report(buf, 8192)
tot = 0
for line in buf.value.decode().strip().splitlines():
    name, k, ms = line.split()
    print(f"{name:28s} {int(k):5d} launches  {1e3 * float(ms) / int(k):8.2f} us avg  {1e3 * float(ms) / n:9.1f} us / iteration")
    tot += 1e3 * float(ms) / n
print("sum per iteration (us):", tot)
