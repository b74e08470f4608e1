"""Development check (GPU box): product vs GPU oracle on one config.  Not part of the test suite."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from monogs_b200 import scene as S, rasterizer as RZ, _lib as L
from oracle import ref_gpu

name = sys.argv[1] if len(sys.argv) > 1 else "C1"
P = int(sys.argv[2]) if len(sys.argv) > 2 else None
sc, cam = S.make_config(name, P=P)
dev = "cuda"
cam = S.Camera(*[x.to(dev) if isinstance(x, torch.Tensor) else x for x in cam])
sc = S.Scene(*[x.to(dev) if isinstance(x, torch.Tensor) else x for x in sc])
bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
rs = RZ.GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, bg, 1.0, cam.viewmatrix, cam.projmatrix,
                                      cam.projmatrix_raw, sc.sh_degree, cam.campos, False, False)
gc, gd = S.output_grads(cam.W, cam.H)
gc, gd = gc.to(dev), gd.to(dev)

def rel(a, b):
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))

color, radii, depth, opacity, n_touched, st = RZ.raster_forward(rs, sc.means3D, sc.opacities, sc.shs, None, sc.scales, sc.rotations, None)
torch.cuda.synchronize()
fw = ref_gpu.forward(cam, bg, sc.means3D, sc.opacities, shs=sc.shs, scales=sc.scales, rotations=sc.rotations, sh_degree=sc.sh_degree)
torch.cuda.synchronize()
print("R mine/ref", st.num_rendered, fw["R"], "P_vis", int((radii > 0).sum()))
print("radii mismatches", int((radii != fw["radii"]).sum()))
print("color rel", rel(color, fw["color"]), "depth rel", rel(depth, fw["depth"]), "opacity rel", rel(opacity, fw["opacity"]))
print("n_touched mismatches", int((n_touched != fw["n_touched"]).sum()))
lib = L.lib()
Lay_keys = None
# sorted keys / list: read from my workspace via layout offsets exposed through a tiny helper
keys_ref, list_ref = ref_gpu.binning(fw)
ranges_ref, finalT_ref, ncontrib_ref = ref_gpu.image_state(cam)
import ctypes as C
off = (C.c_size_t * 32)()
lib.mgs_debug_layout.restype = C.c_int
lib.mgs_debug_layout(cam.W * 0 + sc.means3D.size(0), cam.W, cam.H, C.c_int64(st.num_rendered), off)
names = ["header","depths","means2D","conic_opacity","rgb4","tiles_touched","point_offsets","final_T","n_contrib","ranges","records","keys_in","keys_out","vals_in","vals_out"]
o = dict(zip(names, list(off)))
ws = st.workspace
R = st.num_rendered
list_mine = ws[o["vals_out"]:o["vals_out"] + 4 * R].view(torch.int32)
Pn = sc.means3D.size(0)
keys_mine = (ws[o["keys_out"]:o["keys_out"] + 4 * R].view(torch.int32).long() << 32) | (ws[o["depths"]:o["depths"] + 4 * Pn].view(torch.int32)[list_mine.long()].long() & 0xFFFFFFFF)
tiles = ((cam.W + 15) // 16) * ((cam.H + 15) // 16)
ranges_mine = ws[o["ranges"]:o["ranges"] + 8 * tiles].view(torch.int32).view(tiles, 2)
print("keys equal", bool((keys_mine == keys_ref).all()), "list equal", bool((list_mine == list_ref).all()),
      "ranges equal", bool((ranges_mine == ranges_ref).all()))
nc_mine = ws[o["n_contrib"]:o["n_contrib"] + 4 * cam.W * cam.H].view(torch.int32).view(cam.H, cam.W)
print("n_contrib mismatches", int((nc_mine != ncontrib_ref).sum()))

g = RZ.raster_backward(rs, st, L.GRAD_ALL, sc.means3D, sc.shs, None, sc.scales, sc.rotations, None, radii, gc, gd, want_tau_per_gaussian=True)
torch.cuda.synchronize()
gr = ref_gpu.backward(cam, bg, fw, sc.means3D, gc, gd, shs=sc.shs, scales=sc.scales, rotations=sc.rotations, sh_degree=sc.sh_degree)
torch.cuda.synchronize()
print("grad means3D", rel(g["means3D"], gr["means3D"]), "means2D", rel(g["means2D"], gr["means2D"]),
      "opacity", rel(g["opacity"], gr["opacity"]), "sh", rel(g["color"], gr["sh"]),
      "scales", rel(g["scales"], gr["scales"]), "rot", rel(g["rotations"], gr["rotations"]))
print("tau_pg", rel(g["tau_pg"], gr["tau"]), "tau sum mine", g["tau"].tolist(), "ref", gr["tau"].sum(0).tolist())

def timeit(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
t_f = timeit(lambda: RZ.raster_forward(rs, sc.means3D, sc.opacities, sc.shs, None, sc.scales, sc.rotations, None))
t_b = timeit(lambda: RZ.raster_backward(rs, st, L.GRAD_ALL, sc.means3D, sc.shs, None, sc.scales, sc.rotations, None, radii, gc, gd))
t_rf = timeit(lambda: ref_gpu.forward(cam, bg, sc.means3D, sc.opacities, shs=sc.shs, scales=sc.scales, rotations=sc.rotations, sh_degree=sc.sh_degree))
t_rb = timeit(lambda: ref_gpu.backward(cam, bg, fw, sc.means3D, gc, gd, shs=sc.shs, scales=sc.scales, rotations=sc.rotations, sh_degree=sc.sh_degree))
print(f"ms: mine fwd {t_f:.3f} bwd {t_b:.3f} | ref fwd {t_rf:.3f} bwd {t_rb:.3f}")
