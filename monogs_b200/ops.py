"""Host-side wrappers for the fused loss, fused Adam, distCUDA2 and the on-device pose update.

Each is a thin ctypes call into libmonogs_b200.so; torch only provides memory and the current stream.
Reference behaviour being mirrored (MonoGS tree): utils/slam_utils.py:58-113,188-253 (losses),
gaussian_splatting/scene/gaussian_model.py:247-296 + torch.optim.Adam (optimizer),
submodules/simple-knn/spatial.cu:15-26 (distCUDA2), utils/pose_utils.py:88-98 (update_pose).
"""
import ctypes as C

import torch

from . import _lib as L
from .rasterizer import _f32c, _stream


def distCUDA2(points: torch.Tensor) -> torch.Tensor:
    """float[P]: mean squared distance of every point to its three nearest other points."""
    pts = _f32c(points)
    P = points.size(0)
    out = torch.zeros(P, device=points.device, dtype=torch.float32)
    if P == 0:
        return out
    lib = L.lib()
    ws = torch.empty(lib.mgs_knn_workspace_bytes(P), device=points.device, dtype=torch.uint8)
    L.check(lib.mgs_knn_dist2(P, L.ptr(pts), L.ptr(out), L.ptr(ws), ws.numel(), _stream(points.device)), "mgs_knn_dist2")
    return out


class FusedLoss(torch.autograd.Function):
    """loss(image, depth, exposure_a, exposure_b) with the gradient images produced in the same sweep."""

    @staticmethod
    def forward(ctx, image, depth, exposure_a, exposure_b, opacity, gt_image, gt_depth, grad_mask, mode,
                rgb_boundary_threshold, alpha, huber_delta):
        lib = L.lib()
        dev = image.device
        _, H, W = image.shape
        use_depth = gt_depth is not None and depth is not None and mode != L.LOSS_TRACK_HUBER
        a = L.MgsLossArgs()
        a.mode, a.H, a.W, a.use_depth = mode, H, W, int(use_depth)
        a.rgb_boundary_threshold, a.alpha, a.huber_delta, a.loss_scale = rgb_boundary_threshold, alpha, huber_delta, 1.0
        ts = dict(image=_f32c(image), depth=_f32c(depth) if use_depth else None, opacity=_f32c(opacity),
                  gt_image=_f32c(gt_image), gt_depth=_f32c(gt_depth) if use_depth else None,
                  grad_mask=_f32c(grad_mask.float()) if grad_mask is not None else None,
                  exposure_a=_f32c(exposure_a), exposure_b=_f32c(exposure_b))
        for k, t in ts.items():
            setattr(a, k, t.data_ptr() if t is not None else None)
        loss = torch.empty(1, device=dev)
        dab = torch.empty(2, device=dev)
        g_img = torch.empty(3, H, W, device=dev)
        g_dep = torch.empty(1, H, W, device=dev)
        scratch = torch.empty(lib.mgs_loss_scratch_bytes(H, W), device=dev, dtype=torch.uint8)
        L.check(lib.mgs_loss_fwd_bwd(C.byref(a), L.ptr(loss), L.ptr(dab), L.ptr(g_img), L.ptr(g_dep), L.ptr(scratch),
                                     scratch.numel(), _stream(dev)), "mgs_loss_fwd_bwd")
        ctx.save_for_backward(g_img, g_dep, dab)
        ctx.has = (depth is not None, exposure_a is not None, exposure_b is not None)
        return loss.squeeze(0)

    @staticmethod
    def backward(ctx, g):
        g_img, g_dep, dab = ctx.saved_tensors
        has_d, has_a, has_b = ctx.has
        return (g * g_img, g * g_dep if has_d else None, (g * dab[0]).view(1) if has_a else None,
                (g * dab[1]).view(1) if has_b else None, None, None, None, None, None, None, None, None)


def tracking_loss(image, depth, opacity, gt_image, exposure_a, exposure_b, grad_mask=None, gt_depth=None,
                  rgb_boundary_threshold=0.01, alpha=0.95):
    """get_loss_tracking (slam_utils.py:83-113): mono when gt_depth is None, else rgbd."""
    return FusedLoss.apply(image, depth, exposure_a, exposure_b, opacity, gt_image, gt_depth, grad_mask,
                           L.LOSS_TRACK_L1, rgb_boundary_threshold, alpha, 0.01)


def tracking_loss_huber(image, opacity, gt_image, exposure_a, exposure_b, rgb_boundary_threshold=0.01, delta=0.01):
    """‖Huber_delta(opacity * (image_ab*m - gt*m))‖₂ — the fork's first-order tracking objective
    (slam_frontend.py:506,596-598; slam_utils.py:58-75,188-205)."""
    return FusedLoss.apply(image, None, exposure_a, exposure_b, opacity, gt_image, None, None, L.LOSS_TRACK_HUBER,
                           rgb_boundary_threshold, 0.95, delta)


def mapping_loss(image, depth, gt_image, exposure_a=None, exposure_b=None, gt_depth=None,
                 rgb_boundary_threshold=0.01, alpha=0.95):
    """get_loss_mapping (slam_utils.py:224-253); exposure None = initialization (image_ab = image)."""
    return FusedLoss.apply(image, depth, exposure_a, exposure_b, None, gt_image, gt_depth, None, L.LOSS_MAP_L1,
                           rgb_boundary_threshold, alpha, 0.01)


class FusedAdam:
    """torch.optim.Adam-compatible subset (param_groups with per-group lr, betas, eps; no weight decay /
    amsgrad) executed as one multi-tensor kernel launch per step.

    capturable=True: the step number and the learning rates live in device memory (mgs_adam_step_device, the
    lr_device field of mgs_adam_group), so step() can be captured into a CUDA graph once and replayed while
    set_lr() keeps a schedule going (GaussianModel.update_learning_rate changes the xyz rate every iteration,
    gaussian_model.py:298-312).  All parameters then share one step counter."""

    def __init__(self, params, lr=0.0, betas=(0.9, 0.999), eps=1e-8, capturable=False):
        groups = list(params)
        if groups and not isinstance(groups[0], dict):
            groups = [{"params": groups}]
        self.param_groups = []
        for g in groups:
            g = dict(g)
            g["params"] = list(g["params"]) if not isinstance(g["params"], torch.Tensor) else [g["params"]]
            g.setdefault("lr", lr); g.setdefault("betas", betas); g.setdefault("eps", eps)
            self.param_groups.append(g)
        self.state = {}
        self.betas, self.eps = betas, eps
        self.capturable = capturable
        self._step_dev = self._lr_dev = None

    def zero_grad(self, set_to_none=True):
        for g in self.param_groups:
            for p in g["params"]:
                if p.grad is not None:
                    if set_to_none: p.grad = None
                    else: p.grad.zero_()

    def set_lr(self, group_index: int, lr: float):
        """Install a new learning rate for one group (host value and, when capturable, its device copy)."""
        self.param_groups[group_index]["lr"] = float(lr)
        if self._lr_dev is not None:
            self._lr_dev[group_index].fill_(float(lr))

    def device_step(self):
        """The device-side step counter of a capturable optimiser (i32[1]); None before the first step."""
        return self._step_dev

    def set_step(self, n: int):
        for g in self.param_groups:
            for p in g["params"]:
                if p in self.state:
                    self.state[p]["step"] = int(n)
        if self._step_dev is not None:
            self._step_dev.fill_(int(n))
        self._pending_step = int(n)

    def _state_of(self, p):
        st = self.state.get(p)
        if st is None:
            st = self.state[p] = dict(step=0, exp_avg=torch.zeros_like(p, memory_format=torch.contiguous_format),
                                      exp_avg_sq=torch.zeros_like(p, memory_format=torch.contiguous_format))
        return st

    @torch.no_grad()
    def step(self):
        lib = L.lib()
        if self.capturable:
            items = [(gi, g, p) for gi, g in enumerate(self.param_groups) for p in g["params"] if p.grad is not None]
            if not items:
                return
            dev = items[0][2].device
            if self._step_dev is None:      # created outside any capture: the first step of a capturable optimiser is eager
                start = max([self._state_of(p)["step"] for _, _, p in items] + [getattr(self, "_pending_step", 0)])
                self._step_dev = torch.full((1,), start, device=dev, dtype=torch.int32)
                self._lr_dev = torch.tensor([g["lr"] for g in self.param_groups], device=dev, dtype=torch.float32)
            arr = (L.MgsAdamGroup * len(items))()
            for i, (gi, g, p) in enumerate(items):
                st = self._state_of(p)
                grad = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
                assert p.is_contiguous() and p.dtype == torch.float32
                arr[i].param, arr[i].grad = p.data_ptr(), grad.data_ptr()
                arr[i].exp_avg, arr[i].exp_avg_sq = st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr()
                arr[i].numel, arr[i].lr = p.numel(), float(g["lr"])
                arr[i].lr_device = self._lr_dev[gi:gi + 1].data_ptr()
            b1, b2 = self.param_groups[0]["betas"]
            L.check(lib.mgs_adam_step_device(arr, len(items), L.ptr(self._step_dev), b1, b2, self.param_groups[0]["eps"],
                                             _stream(dev)), "mgs_adam_step_device")
            return
        by_step = {}
        for g in self.param_groups:
            for p in g["params"]:
                if p.grad is None:
                    continue
                st = self._state_of(p)
                st["step"] += 1
                grad = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
                by_step.setdefault((st["step"], g["betas"], g["eps"]), []).append((p, grad, st, g["lr"]))
        for (step, betas, eps), items in by_step.items():
            arr = (L.MgsAdamGroup * len(items))()
            for i, (p, grad, st, lr) in enumerate(items):
                assert p.is_contiguous() and p.dtype == torch.float32
                arr[i].param, arr[i].grad = p.data_ptr(), grad.data_ptr()
                arr[i].exp_avg, arr[i].exp_avg_sq = st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr()
                arr[i].numel, arr[i].lr = p.numel(), float(lr)
            L.check(lib.mgs_adam_step(arr, len(items), step, betas[0], betas[1], eps, _stream(items[0][0].device)), "mgs_adam_step")


def pose_update(T_w2c, rho, theta, proj_raw=None, want_matrices=False, converged_threshold=1e-4):
    """In-place T <- Exp(rho,theta) T on the device; returns (converged i32[1] tensor, viewmatrix, projmatrix)."""
    lib = L.lib()
    dev = T_w2c.device
    assert T_w2c.is_contiguous() and T_w2c.dtype == torch.float32
    conv = torch.empty(1, device=dev, dtype=torch.int32)
    view = torch.empty(4, 4, device=dev) if want_matrices else None
    proj = torch.empty(4, 4, device=dev) if (want_matrices and proj_raw is not None) else None
    L.check(lib.mgs_pose_update(L.ptr(T_w2c), L.ptr(rho), L.ptr(theta), L.ptr(_f32c(proj_raw)), L.ptr(view), L.ptr(proj),
                                L.ptr(conv), converged_threshold, _stream(dev)), "mgs_pose_update")
    return conv, view, proj
