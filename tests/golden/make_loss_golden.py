"""Generates tests/golden/loss_*.npz by running the REFERENCE'S OWN Python code for the rows around the rasterizer
 — imported unchanged from /root/reference, on the CPU, in THIS container:

    utils/slam_utils.py   get_loss_tracking (rgb / rgbd), get_loss_tracking_per_pixel + HuberLoss (the fork's first-order
                          objective, slam_frontend.py:506,596-598), the second-order sketched objective with
                          ApplyExposure's sketched exposure columns (slam_utils.py:115-185, slam_frontend.py:637-665),
                          get_loss_mapping (rgb / rgbd / initialization), image_gradient, image_gradient_mask,
                          get_median_depth
    utils/camera_utils.py Camera.compute_grad_mask (global median and the Replica per-patch variant)
    utils/pose_utils.py   SE3_exp / update_pose arithmetic (the pure-torch restatement the fork ships beside lietorch)

    python tests/golden/make_loss_golden.py            # writes next to this file

The only shims are environmental: utils.configs.cuda_device is set to "cpu" before utils.slam_utils binds it (the
module has no other device dependence), torch.cuda.synchronize is a no-op (ApplyExposure.backward calls it for its
timers) and an empty `lietorch` module satisfies pose_utils' import (lietorch 0.2 is not installed; the pure-torch SE3_exp
that is exported here does not touch it).  /root/reference does not travel to the GPU box, so the vectors are committed and tests/test_oracle_cpu.py
checks the repo's restatements (oracle/golden.py: loss_ref, sketch_loss_ref, image_gradient*, grad_masks_ref,
median_depth_ref, se3_exp_ref) against them; the GPU tests then compare the CUDA ops with those restatements.
"""
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REF)
import utils.configs as ref_cfg          # noqa: E402
ref_cfg.cuda_device = "cpu"
torch.cuda.synchronize = lambda *a, **k: None
sys.modules.setdefault("lietorch", types.ModuleType("lietorch"))   # pose_utils imports it; SE3_exp below does not use it
import utils.slam_utils as SU            # noqa: E402  (the reference module, unchanged)
import utils.camera_utils as CU          # noqa: E402
import utils.pose_utils as PU            # noqa: E402

torch.manual_seed(0)
H, W = 40, 56


def images(seed):
    g = torch.Generator().manual_seed(seed)
    image = torch.rand(3, H, W, generator=g)
    gt = torch.rand(3, H, W, generator=g)
    gt[:, :4, :] = 0.0                              # below rgb_boundary_threshold -> masked out
    depth = torch.rand(1, H, W, generator=g) * 3
    gt_depth = torch.rand(1, H, W, generator=g) * 3
    gt_depth[:, :, :5] = 0.0
    opacity = torch.rand(1, H, W, generator=g)
    opacity[:, 20:, :] = opacity[:, 20:, :] * 0.05 + 0.95
    return image, gt, depth, gt_depth, opacity


def viewpoint(gt, gt_depth, a, b, config):
    vp = types.SimpleNamespace()
    vp.original_image, vp.device = gt, "cpu"
    vp.depth = gt_depth[0].numpy()
    vp.exposure_a, vp.exposure_b, vp.exposure_eps = a, b, 1e-8       # camera_utils.py:55-61
    CU.Camera.compute_grad_mask(vp, config)                            # grad_mask, rgb_pixel_mask(_mapping), gt_depth
    return vp


def cfg(mono, dataset="tum"):
    return {"Training": {"monocular": mono, "rgb_boundary_threshold": 0.01, "edge_threshold": 1.1, "alpha": 0.95},
            "Dataset": {"type": dataset}}


out = {}
image, gt, depth, gt_depth, opacity = images(0)
out.update(image=image, gt=gt, depth=depth, gt_depth=gt_depth, opacity=opacity)
a0, b0 = torch.tensor([-0.7]), torch.tensor([0.05])
out.update(a=a0, b=b0)


def leafs():
    return [t.clone().requires_grad_(True) for t in (image, depth, a0, b0)]


# ---- a12: tracking / mapping losses and their gradients ----
for name, mono in (("track_rgb", True), ("track_rgbd", False)):
    im, dp, a, b = leafs()
    c = cfg(mono)
    vp = viewpoint(gt, gt_depth, a, b, c)
    loss = SU.get_loss_tracking(c, im, dp, opacity, vp)
    loss.backward()
    out[name + "_loss"] = loss.detach()
    out[name + "_g_image"], out[name + "_g_a"], out[name + "_g_b"] = im.grad, a.grad, b.grad
    if not mono:
        out[name + "_g_depth"] = dp.grad
    out["grad_mask"] = vp.grad_mask.float()

im, dp, a, b = leafs()
c = cfg(True)
vp = viewpoint(gt, gt_depth, a, b, c)
r = SU.get_loss_tracking_per_pixel(c, im, dp, opacity, vp)                 # slam_frontend.py:506
out["huber_scalar_l1"] = torch.norm(r.flatten(), p=1).detach()              # :510 (bookkeeping)
loss = torch.norm(SU.HuberLoss.apply(r, 0.01).flatten(), p=2)               # :596-598
loss.backward()
out["huber_loss"], out["huber_g_image"], out["huber_g_a"], out["huber_g_b"] = loss.detach(), im.grad, a.grad, b.grad

for name, mono, init in (("map_rgb", True, False), ("map_rgbd", False, False), ("map_init", True, True)):
    im, dp, a, b = leafs()
    c = cfg(mono)
    vp = viewpoint(gt, gt_depth, a, b, c)
    loss = SU.get_loss_mapping(c, im, dp, vp, opacity, initialization=init)
    loss.backward()
    out[name + "_loss"], out[name + "_g_image"] = loss.detach(), im.grad
    if not init:
        out[name + "_g_a"], out[name + "_g_b"] = a.grad, b.grad
    if not mono:
        out[name + "_g_depth"] = dp.grad

# ---- a12 / a16: the second-order sketched objective (slam_frontend.py:269-338, 637-665) on a tiny sketch ----
repeat_dim, stack_dim, sketch_dim = 1, 4, 8
m, d = H * W, stack_dim * sketch_dim
chunk = m // d
g = torch.Generator().manual_seed(11)
perm = torch.randperm(m, generator=g, dtype=torch.int64)[:chunk * d].to(torch.int32)
rows = (perm // W).view(1, stack_dim, sketch_dim, chunk).long()
cols = (perm % W).view(1, stack_dim, sketch_dim, chunk).long()
sketch_indices = torch.full((repeat_dim, stack_dim, H, W), -1, dtype=torch.int32)
for j in range(stack_dim):
    for i in range(sketch_dim):
        sketch_indices[0, j, rows[0, j, i], cols[0, j, i]] = i
rand_weights = (torch.randint(0, 2, (repeat_dim, H, W), generator=g).float() * 2 - 1)
sketch_dexposure = torch.empty(stack_dim, sketch_dim, 2, requires_grad=True)
im, dp, a, b = leafs()
c = cfg(True)
vp = viewpoint(gt, gt_depth, a, b, c)
fsa = {"sketch_mode": 1, "sketch_dim": sketch_dim, "stack_dim": stack_dim, "rand_indices": (rows, cols),
       "sketch_dexposure": sketch_dexposure}
r = SU.get_loss_tracking_per_pixel(c, im, dp, opacity, vp, forward_sketch_args=fsa)
li = SU.HuberLoss.apply(r, 0.01).sum(dim=0) / (m / d)                       # slam_frontend.py:641-642
weighted = li * rand_weights                                                # :648
Sf = weighted[torch.arange(repeat_dim).view(-1, 1, 1, 1), rows, cols].sum(dim=-1)   # :651
weighted[0].backward(gradient=torch.ones_like(weighted[0]), retain_graph=True)        # :663
out.update(sk_indices=sketch_indices, sk_weights=rand_weights, sk_Sf=Sf.detach(), sk_g_image=im.grad,
           sk_dexposure=sketch_dexposure.grad, sk_g_a=a.grad, sk_g_b=b.grad, sk_rows=rows.int(), sk_cols=cols.int())

# ---- f4: frame masks and medians ----
gi = torch.Generator().manual_seed(5)
frame = torch.rand(3, 64, 96, generator=gi)
frame[:, :10, :30] = 0.0
gray = frame.mean(dim=0, keepdim=True)
gv, gh = SU.image_gradient(gray)
mv, mh = SU.image_gradient_mask(gray)
out.update(frame=frame, frame_gv=gv, frame_gh=gh, frame_mv=mv, frame_mh=mh)
for ds in ("tum", "replica"):
    vp = types.SimpleNamespace(original_image=frame, device="cpu", depth=None)
    CU.Camera.compute_grad_mask(vp, cfg(True, ds))
    out[f"frame_grad_mask_{ds}"] = vp.grad_mask.float()
    out[f"frame_rgb_pixel_mask_{ds}"] = vp.rgb_pixel_mask.float()
    out[f"frame_rgb_pixel_mask_mapping_{ds}"] = vp.rgb_pixel_mask_mapping.float()
dep = torch.rand(1, 64, 96, generator=gi) * 4 - 0.5
opa = torch.rand(1, 64, 96, generator=gi) * 0.1 + 0.92
msk = torch.rand(1, 64, 96, generator=gi) > 0.3
out.update(med_depth=dep, med_opacity=opa, med_mask=msk,
           med_plain=SU.get_median_depth(dep, opa), med_masked=SU.get_median_depth(dep, opa, msk))

# ---- f2: SE3 exponential (utils/pose_utils.py:26-74) ----
taus = torch.tensor([[0.01, -0.005, 0.008, 0.004, -0.003, 0.002], [1e-7, 0, 0, 0, 1e-7, 0],
                     [0.2, 0.1, -0.3, 0.5, -0.4, 0.3], [0.0, 0.0, 0.0, 0.0, 0.0, 0.0]])
out["se3_tau"] = taus
out["se3_exp"] = torch.stack([PU.SE3_exp(t) for t in taus])

np.savez_compressed(os.path.join(HERE, "loss_ref.npz"),
                    **{k: (v.detach().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)) for k, v in out.items()})
print("wrote", os.path.join(HERE, "loss_ref.npz"), len(out), "arrays")
