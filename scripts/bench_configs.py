"""Timing sweep over BASELINE.json's configs C1 / C2 / C3 / C5 on one B200 (the bench line itself stays on C2):
rasterizer fwd+bwd through the C-ABI vs the unmodified reference CUDA build on identical inputs, the fused tracking
iteration (C2 mono, C3 rgbd) and the MapEngine mapping iteration (C3: 10-frame rgbd window, C5: one view + Adam).
Prints one JSON object; the numbers are copied into BASELINE.md §3.  Usage: python scripts/bench_configs.py [C1 C2 ...]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from monogs_b200 import scene as S, rasterizer as RZ, _lib as L
from monogs_b200.tracking import PoseTracker
from monogs_b200.mapping import MapEngine, MapView
from oracle import ref_gpu

dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, device=dev, dtype=torch.uint8)


def timed(fn, steps, warmup=3):
    for _ in range(warmup):
        flush.zero_(); fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for a, b in ev:
        flush.zero_(); a.record(); fn(); b.record()
    torch.cuda.synchronize()
    ms = sorted(a.elapsed_time(b) for a, b in ev)
    return ms[len(ms) // 2]


def to_dev(nt):
    return type(nt)(*[x.to(dev) if isinstance(x, torch.Tensor) else x for x in nt])


def run(cfg):
    sc, cam = S.make_config(cfg)
    scd, camd = to_dev(sc), to_dev(cam)
    bg = torch.zeros(3, device=dev)
    gc, gd = S.output_grads(cam.W, cam.H)
    gc, gd = gc.to(dev), gd.to(dev)
    rs = RZ.GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, bg, 1.0, camd.viewmatrix, camd.projmatrix,
                                          camd.projmatrix_raw, sc.sh_degree, camd.campos, False, False)
    out = {"config": cfg, "gaussians": sc.means3D.size(0), "image": [cam.W, cam.H], "sh_degree": sc.sh_degree}
    steps = 20 if cfg != "C5" else 8
    st = {}

    def ours():
        color, radii, depth, opacity, n_touched, s_ = RZ.raster_forward(rs, scd.means3D, scd.opacities, scd.shs, None, scd.scales,
                                                                        scd.rotations, None)
        st["s"], st["radii"] = s_, radii
        RZ.raster_backward(rs, s_, L.GRAD_ALL, scd.means3D, scd.shs, None, scd.scales, scd.rotations, None, radii, gc, gd)
    out["raster_fwd_bwd_ms"] = timed(ours, steps)
    out["tile_instances"], out["visible"] = st["s"].num_rendered, int((st["radii"] > 0).sum())
    if ref_gpu.available():
        def ref():
            fw = ref_gpu.forward(camd, bg, scd.means3D, scd.opacities, shs=scd.shs, scales=scd.scales, rotations=scd.rotations,
                                 sh_degree=sc.sh_degree)
            ref_gpu.backward(camd, bg, fw, scd.means3D, gc, gd, shs=scd.shs, scales=scd.scales, rotations=scd.rotations,
                             sh_degree=sc.sh_degree)
        out["ref_cuda_raster_fwd_bwd_ms"] = timed(ref, max(steps // 2, 4))
        out["raster_speedup"] = out["ref_cuda_raster_fwd_bwd_ms"] / out["raster_fwd_bwd_ms"]
    # target frames: the scene at a perturbed pose
    T_gt = S.se3_exp(torch.tensor(S.GT_TAU)).float()
    cam_gt = to_dev(S.make_camera(cam.W, cam.H, cam.fx, cam.fy, cam.cx, cam.cy, T_gt))
    rs_gt = rs._replace(viewmatrix=cam_gt.viewmatrix, projmatrix=cam_gt.projmatrix, campos=cam_gt.campos)
    gt, _, gt_depth, _, _, _ = RZ.raster_forward(rs_gt, scd.means3D, scd.opacities, scd.shs, None, scd.scales, scd.rotations, None)
    if sc.sh_degree == 0:
        rgbd = cfg == "C3"
        tr = PoseTracker(scd.means3D, scd.opacities, scd.shs, scd.scales, scd.rotations, 0, cam, bg, torch.eye(4),
                         mode=L.LOSS_TRACK_L1 if rgbd else L.LOSS_TRACK_HUBER, use_depth=rgbd)
        tr.set_frame(gt, gt_depth=gt_depth if rgbd else None)
        out["tracking_iter_ms"] = timed(tr.step, 50, warmup=10)
        out["tracking_loss"] = "rgbd L1 (alpha 0.95)" if rgbd else "mono Huber+L2"
        out["tracking_overflow"] = tr.check_capacity()[1]
        del tr
    # mapping on the raw parameters
    n_views = {"C1": 8, "C2": 8, "C3": 10, "C5": 1}[cfg]
    oc = scd.opacities.clamp(1e-4, 1 - 1e-4)
    raw = {"xyz": scd.means3D.clone(), "f_dc": scd.shs[:, :1].contiguous(), "opacity": torch.log(oc / (1 - oc)),
           "scaling": torch.log(scd.scales), "rotation": scd.rotations.clone()}
    lrs = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 2.5e-3 / 20, "opacity": 0.05, "scaling": 1e-3, "rotation": 1e-3}
    if scd.shs.size(1) > 1:
        raw["f_rest"] = scd.shs[:, 1:].contiguous()
    eng = MapEngine(raw, lrs, sh_degree=sc.sh_degree, W=cam.W, H=cam.H, capacity_headroom=1.5)
    views = []
    for i in range(n_views):
        tau = torch.tensor(S.GT_TAU) * (i - n_views / 2) * 0.5
        c = to_dev(S.make_camera(cam.W, cam.H, cam.fx, cam.fy, cam.cx, cam.cy, S.se3_exp(tau).float()))
        v = rs._replace(viewmatrix=c.viewmatrix, projmatrix=c.projmatrix, campos=c.campos)
        views.append(MapView(v, (gt * 0.9 + 0.05).contiguous(), gt_depth.contiguous() if cfg == "C3" else None))
    out["mapping_iter_ms"] = timed(lambda: eng.step(views), 10 if cfg != "C5" else 5, warmup=2)
    out["mapping_views"], out["mapping_overflow"] = n_views, eng.check_capacity()[1]
    out["mapping_what"] = f"{n_views} view(s): render + mapping loss{' (rgbd)' if cfg == 'C3' else ''} + full backward, regulariser, fused Adam"
    del eng
    torch.cuda.empty_cache()
    return out


if __name__ == "__main__":
    cfgs = sys.argv[1:] or ["C1", "C2", "C3", "C5"]
    res = [run(c) for c in cfgs]
    print(json.dumps({"results": res, "gpu": torch.cuda.get_device_name(0)}))
